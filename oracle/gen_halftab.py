#!/usr/bin/env python3
"""oracle/gen_halftab.py -- TEST INFRASTRUCTURE ONLY.

Runs the reference's own table generator (src/cprocess/genhalf.py, Python-2
`print` statements; SConstruct:109 pipes its stdout into halftab.c) under
python3 and writes the generated C to oracle/_ref/halftab.c (git-ignored
build output).  The generator is executed from where it lies under the
reference tree; only its *output* lands in this repo's ignored build dir.
"""
import contextlib
import io
import re
import sys


def main(genhalf_path: str, out_path: str) -> None:
    src = open(genhalf_path, "r", encoding="utf-8").read()
    lines = []
    for line in src.splitlines():
        m = re.match(r"^(\s*)print\s*$", line)
        if m:
            lines.append(f"{m.group(1)}print()")
            continue
        m = re.match(r"^(\s*)print (.*)$", line)
        if m:
            # py2 `print a, b` joins with single spaces, same as print(a, b)
            lines.append(f"{m.group(1)}print({m.group(2)})")
            continue
        lines.append(line)
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        exec(compile("\n".join(lines), genhalf_path, "exec"), {"__name__": "__main__"})
    with open(out_path, "w", encoding="utf-8") as f:
        f.write(buf.getvalue())


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
