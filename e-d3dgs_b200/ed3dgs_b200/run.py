"""python -m ed3dgs_b200.run <script.py> [args...]: runs an unmodified script of the reference checkout (train.py,
render.py, mesh_extract_tetrahedra.py) with the drop-ins installed (dropin.install) -- no source edits."""
import os
import runpy
import sys


def main(argv=None):
    argv = list(sys.argv[1:] if argv is None else argv)
    if not argv:
        raise SystemExit("usage: python -m ed3dgs_b200.run <script.py> [args...]")
    from . import dropin
    dropin.install()
    script = argv[0]
    sys.argv = argv
    sys.path.insert(0, os.path.dirname(os.path.abspath(script)))   # what `python script.py` does
    runpy.run_path(script, run_name="__main__")


if __name__ == "__main__":
    main()
