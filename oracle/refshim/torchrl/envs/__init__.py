"""TEST INFRASTRUCTURE ONLY — stub of torchrl.envs.EnvBase (see torchrl/__init__)."""
import torch


class EnvBase:
    batch_locked = False

    def __init__(self, device="cpu", batch_size=None, run_type_checks=False,
                 allow_done_after_reset=False, **kw):
        self.device = torch.device(device) if device is not None else torch.device("cpu")
        self.batch_size = torch.Size(batch_size) if batch_size is not None else torch.Size([])
        self.run_type_checks = run_type_checks
        self.allow_done_after_reset = allow_done_after_reset

    def reset(self, td=None, **kw):
        out = self._reset(td, **kw)
        bs = tuple(out.batch_size)
        for key in ("done", "terminated"):
            if key not in out:
                out.set(key, torch.zeros(*bs, 1, dtype=torch.bool, device=out.device))
        return out

    def set_seed(self, seed=None, static_seed=False):
        self._set_seed(seed)
        return seed

    def to(self, device):
        self.device = torch.device(device)
        return self

    def _assert_tensordict_shape(self, td):
        pass

    def _step_proc_data(self, td):
        return td
