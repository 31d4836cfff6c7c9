"""TEST INFRASTRUCTURE ONLY — inert stand-in for `torchrl` (not installable here).
Only the bookkeeping the reference's RL4COEnvBase needs: reset -> _reset + default
`done`, set_seed -> _set_seed.  No arithmetic.  See SURVEY.md Appendix A."""
__version__ = "0.7.0"
