"""TEST INFRASTRUCTURE ONLY -- see data_readers/__init__.py."""


def compute_distance_matrix_flow(*args, **kwargs):
    raise NotImplementedError("training-only helper; not part of the hot path")


def compute_distance_matrix_flow2(*args, **kwargs):
    raise NotImplementedError("training-only helper; not part of the hot path")
