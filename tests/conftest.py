import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def built():
    """Make sure the native artefacts exist (the driver normally ran __graft_entry__.build() already)."""
    import __graft_entry__ as g
    need = [os.path.join(ROOT, "atray_b200", "libatray_b200.so"), os.path.join(ROOT, "oracle", "liboracle.so")]
    if not all(os.path.exists(p) for p in need):
        g.build()
    return True


@pytest.fixture(scope="session")
def port(built):
    from oracle import bind
    return bind.port()


@pytest.fixture(scope="session")
def ref(built):
    from oracle import bind
    if not bind.ref_available():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    return bind.ref()


@pytest.fixture(scope="session")
def meshes(built):
    """Procedural meshes (always available) + the reference's own assets when staged under oracle/_ref/assets."""
    from oracle import bind
    from tools import gen_standin_obj
    d = os.environ.get("ATRAY_MESH_CACHE", "/tmp/atray_meshes")
    os.makedirs(d, exist_ok=True)
    m = {
        "knot_small": (gen_standin_obj.ensure(os.path.join(d, "knot_48x32.obj"), 48, 32), 0.0),
        "knot_mid": (gen_standin_obj.ensure(os.path.join(d, "knot_96x64.obj"), 96, 64), 0.0),
    }
    for name in ("Simple", "Cube", "Deer", "Monkey"):
        p = os.path.join(bind.ASSETS, name + ".obj")
        if os.path.exists(p):
            m[name] = (p, 40.0)
    return m


@pytest.fixture(scope="session")
def standin(built):
    from tools import gen_standin_obj
    d = os.environ.get("ATRAY_MESH_CACHE", "/tmp/atray_meshes")
    return gen_standin_obj.ensure(os.path.join(d, "standin_272x256.obj"))
