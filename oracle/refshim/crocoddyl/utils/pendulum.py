"""Import-only stand-in for crocoddyl.utils.pendulum (tests/python/problems.py:15-18)."""


class ActuationModelDoublePendulum:  # pragma: no cover
    def __init__(self, *a, **k):
        raise NotImplementedError("robot models are unavailable in this image")


class CostModelDoublePendulum:  # pragma: no cover
    def __init__(self, *a, **k):
        raise NotImplementedError("robot models are unavailable in this image")
