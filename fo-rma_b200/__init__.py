"""forma_b200 — B200 (sm_100a) drop-in for the per-pixel path-tracing loop of fo-rma's
`src/cpu_ray_tracer`.  The product is libforma_rt.so (csrc/, C ABI in include/forma_rt.h);
`tracer` mirrors the reference's Rust module over it.  Import name: `forma_b200` (the directory
is called fo-rma_b200; forma_b200.py at the repo root aliases it)."""
from . import _lib, tracer  # noqa: F401
from .tracer import TraceModel, create_model, save_image, save_image_mt, srgb_table, update, update_progressive  # noqa: F401

__all__ = ["tracer", "TraceModel", "create_model", "update", "update_progressive", "save_image", "save_image_mt", "srgb_table"]
