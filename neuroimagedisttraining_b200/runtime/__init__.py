from .slab import ParamSlab, SlabLayout  # noqa: F401
