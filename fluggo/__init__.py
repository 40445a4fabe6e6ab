"""The `fluggo` namespace of the reference (fluggo/Canvas), here holding only what the per-pixel frame path replaces:
fluggo.media.process (the CPython extension built from canvas_b200/csrc/py/py_process.c) and fluggo.media.basetypes."""
