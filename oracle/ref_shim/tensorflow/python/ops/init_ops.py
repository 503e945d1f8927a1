class Initializer:
    """Base class of the reference's OrthogonalInit (src/model/backend.py); called as init(shape, dtype)."""
