"""TEST INFRASTRUCTURE ONLY — inert spec classes (see torchrl/__init__)."""


class _Spec:
    def __init__(self, *a, **k):
        self.args, self.kwargs = a, k


class Bounded(_Spec):
    pass


class Unbounded(_Spec):
    pass


class Composite(_Spec):
    pass


BoundedTensorSpec = Bounded
UnboundedContinuousTensorSpec = Unbounded
UnboundedDiscreteTensorSpec = Unbounded
CompositeSpec = Composite
