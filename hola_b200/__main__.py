"""``python -m hola_b200 <command>`` - convenience front end.

    build [--variant release|asan|tsan|static]   compile the native artefacts in-tree
    serve [hola arguments...]                     build if needed, then exec build/hola
    load  [hola-load arguments...]                exec build/hola-load
    coldstart [hola-coldstart arguments...]       exec build/hola-coldstart
    paths                                         print where the artefacts are
"""
from __future__ import annotations

import os
import sys

from . import build, paths


def main(argv) -> int:
    if not argv or argv[0] in ("-h", "--help"):
        print(__doc__)
        return 0
    cmd, rest = argv[0], argv[1:]
    if cmd == "build":
        variant = "release"
        if rest[:1] == ["--variant"] and len(rest) >= 2:
            variant = rest[1]
        p = build(variant, quiet=False)
        print(p.hola)
        return 0
    if cmd == "paths":
        p = build()
        for name in ("hola", "hola_load", "libhola", "unit_tests"):
            print("%-10s %s" % (name, getattr(p, name)))
        return 0
    exe = {"serve": "hola", "load": "hola-load", "coldstart": "hola-coldstart"}.get(cmd)
    if exe is None:
        print("unknown command: %s\n%s" % (cmd, __doc__), file=sys.stderr)
        return 2
    p = build()
    path = str(p.hola.parent / exe)
    os.execv(path, [path, *rest])
    return 127  # not reached


if __name__ == "__main__":
    sys.exit(main(sys.argv[1:]))
