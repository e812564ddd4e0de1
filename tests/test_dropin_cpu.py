"""ed3dgs_b200.install(): the reference's `from scene.deformation import deform_network` (scene/gaussian_model.py:24)
resolves to the drop-in with zero source edits, while the rest of the `scene` package stays the checkout's own."""
import os
import subprocess
import sys
import textwrap

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "e-d3dgs_b200")


def _fake_checkout(tmp_path):
    scene = tmp_path / "scene"
    scene.mkdir()
    # same import order as the reference: scene/__init__.py:17 imports gaussian_model, which imports scene.deformation
    (scene / "__init__.py").write_text("from scene.gaussian_model import GaussianModel\nfrom scene import dataset_readers\n")
    (scene / "deformation.py").write_text("class deform_network:\n    origin = 'reference torch module'\n")
    (scene / "dataset_readers.py").write_text("origin = 'reference'\n")
    (scene / "gaussian_model.py").write_text(textwrap.dedent("""
        from scene.deformation import deform_network
        class GaussianModel:
            net = deform_network
    """))
    (tmp_path / "train.py").write_text(textwrap.dedent("""
        import sys
        from scene import GaussianModel
        import scene.dataset_readers, scene.deformation
        print("NET", GaussianModel.net.__module__, scene.dataset_readers.origin, sys.argv[1:])
    """))
    return tmp_path


def _run(code_or_args, cwd):
    env = dict(os.environ, PYTHONPATH=PKG + os.pathsep + os.environ.get("PYTHONPATH", ""))
    return subprocess.run([sys.executable] + code_or_args, cwd=cwd, env=env, capture_output=True, text=True, timeout=300)


def test_runner_redirects_only_the_deformation_module(tmp_path):
    co = _fake_checkout(tmp_path)
    r = _run(["-m", "ed3dgs_b200.run", "train.py", "-s", "x"], co)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "NET ed3dgs_b200.deformation reference ['-s', 'x']" in r.stdout
    # without the hook the checkout's own module is used
    r = _run(["train.py"], co)
    assert r.returncode == 0 and "NET scene.deformation reference" in r.stdout, r.stderr[-2000:]


def test_install_after_the_fact_replaces_or_refuses(tmp_path):
    co = _fake_checkout(tmp_path)
    code = textwrap.dedent("""
        import sys
        sys.path.insert(0, ".")
        import ed3dgs_b200
        import scene.deformation as before          # nobody else holds its class yet? scene/__init__ does -> refuse
        try:
            ed3dgs_b200.install()
            print("INSTALLED")
        except RuntimeError as e:
            print("REFUSED", "scene.gaussian_model" in str(e))
        ed3dgs_b200.uninstall()
    """)
    r = _run(["-c", code], co)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "REFUSED True" in r.stdout
    code = textwrap.dedent("""
        import sys
        sys.path.insert(0, ".")
        import ed3dgs_b200
        assert ed3dgs_b200.install() and ed3dgs_b200.install()      # idempotent
        import scene
        from ed3dgs_b200.deformation import deform_network
        assert scene.GaussianModel.net is deform_network and sys.modules["scene.deformation"].deform_network is deform_network
        ed3dgs_b200.uninstall()
        assert "scene.deformation" not in sys.modules
        print("OK")
    """)
    r = _run(["-c", code], co)
    assert r.returncode == 0 and "OK" in r.stdout, r.stderr[-2000:]
