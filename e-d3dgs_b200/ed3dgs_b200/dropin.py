"""Zero-edit drop-in for the one import of the hot path that PYTHONPATH alone cannot redirect.

`diff_gaussian_rasterization` and `simple_knn` are top-level packages, so putting e-d3dgs_b200/ in front of
PYTHONPATH is enough for them. The deformation network is `scene.deformation` -- a module INSIDE the reference's own
`scene` package (imported at scene/gaussian_model.py:24) -- and shadowing the whole package would take the dataset
readers with it. install() therefore registers an import hook that answers for that one dotted name and leaves the rest
of `scene` to the reference's files, so train.py / render.py / mesh_extract_tetrahedra.py run unchanged:

    python -m ed3dgs_b200.run train.py -s <scene> --configs arguments/nersemble/default.py -r 2

or, from a sitecustomize.py on the path:  import ed3dgs_b200; ed3dgs_b200.install()
"""
import importlib
import importlib.abc
import importlib.machinery
import sys

# reference module name -> module of this package that carries the same public names
REDIRECTS = {"scene.deformation": "ed3dgs_b200.deformation"}


class _DropInFinder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    def find_spec(self, fullname, path=None, target=None):
        if fullname in REDIRECTS:
            return importlib.machinery.ModuleSpec(fullname, self)
        return None

    def create_module(self, spec):
        return importlib.import_module(REDIRECTS[spec.name])   # the module object itself: one class, one identity

    def exec_module(self, module):
        pass


_finder = None


def install():
    """Idempotent. Must run before the reference's `scene` package is imported; a `scene.deformation` that is already
    loaded is replaced in sys.modules (and on its parent package), but modules that did `from scene.deformation import
    deform_network` earlier keep the class they got -- that case raises, because it would silently run the torch MLP."""
    global _finder
    if _finder is None:
        _finder = _DropInFinder()
        sys.meta_path.insert(0, _finder)
    for name, target in REDIRECTS.items():
        old = sys.modules.get(name)
        mod = importlib.import_module(target)
        if old is not None and old is not mod:
            users = [m.__name__ for m in list(sys.modules.values())
                     if m is not None and m is not old and getattr(m, "deform_network", None) is getattr(old, "deform_network", object())]
            if users:
                raise RuntimeError(f"ed3dgs_b200.install(): {name} was already imported by {users}; call install() first")
            sys.modules[name] = mod
            parent = sys.modules.get(name.rpartition(".")[0])
            if parent is not None:
                setattr(parent, name.rpartition(".")[2], mod)
    return True


def uninstall():
    """Removes the hook and the redirected entries (tests)."""
    global _finder
    if _finder is not None:
        try:
            sys.meta_path.remove(_finder)
        except ValueError:
            pass
        _finder = None
    for name, target in REDIRECTS.items():
        if sys.modules.get(name) is sys.modules.get(target):
            sys.modules.pop(name, None)
            parent = sys.modules.get(name.rpartition(".")[0])
            if parent is not None and getattr(parent, name.rpartition(".")[2], None) is sys.modules.get(target):
                delattr(parent, name.rpartition(".")[2])
