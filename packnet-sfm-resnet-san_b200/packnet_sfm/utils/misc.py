"""Small helpers used on the path (reference: utils/misc.py:55-75)."""


def same_shape(shape1, shape2):
    """True when both shapes have the same length and entries."""
    return len(shape1) == len(shape2) and all(a == b for a, b in zip(shape1, shape2))
