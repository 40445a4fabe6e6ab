"""fluggo.media: `process` is the extension module of this library (PyInit_process, canvas_b200/csrc/py/py_process.c),
`basetypes` the value types its API exchanges."""
