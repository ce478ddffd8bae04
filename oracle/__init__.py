"""ORACLE -- test infrastructure only.

A CPU/torch restatement of the reference's inference hot path, used as the parity
checker by `tests/`, `__graft_entry__.smoke()` and `bench.py`'s `cpu_baseline` /
`--impl reference` legs.  The product package (`video_depth_anything_b200`) never
imports anything from here.
"""
