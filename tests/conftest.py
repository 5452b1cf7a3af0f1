import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

DATA = os.path.join(ROOT, "data")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    from oracle_lib import Oracle

    return Oracle()


# Scenes of the parity tests come from the ORACLE's own loader and builder (oracle/oracle_build.inc), not from the product's: the
# GPU then traverses a tree the code under test did not build, and both sides of every comparison read the same arrays.  That the
# product's builder makes byte-identical arrays is a separate CPU test (tests/test_independent.py).
@pytest.fixture(scope="session")
def teapot(oracle):
    return oracle.build_obj(os.path.join(DATA, "teapot.obj"))


@pytest.fixture(scope="session")
def cube(oracle):
    return oracle.build_obj(os.path.join(DATA, "cube.obj"), quad_mode=1)


@pytest.fixture(scope="session")
def atrium(oracle):
    """The Sponza STAND-IN (minipath_b200/scenes.py: atrium_mesh, 263 k triangles) built by the oracle's builder."""
    from oracle_lib import scenes

    p, n, t = scenes.atrium_mesh()
    return oracle.build_mesh(p, t, n)


@pytest.fixture(scope="session")
def atrium_camera(atrium):
    from minipath_b200 import scenes

    return scenes.atrium_camera(*atrium.get_bounding_box())


@pytest.fixture(scope="session")
def teapot_camera():
    """The camera of benches/render_teapot.rs:12-19 and src/cli.rs:13-20."""
    import minipath_b200 as mp

    return mp.Camera.default().look_at((0.0, 2.0, 10.0), (0.0, 1.5, 0.0), (0.0, 1.0, 0.0)).f_number(4.8).focus_distance(10.0)


@pytest.fixture(scope="session")
def cube_camera():
    """SURVEY §8(d) gate camera for the fan-triangulated cube: pinhole-ish f/64."""
    import minipath_b200 as mp

    return mp.Camera.default().look_at((1.5, 1.2, 2.5), (0.0, 0.0, 0.0), (0.0, 1.0, 0.0)).f_number(64.0)


@pytest.fixture(scope="session", params=["pool", "hybrid_pool", "hybrid_pool_s0", "hybrid", "persist", "octet", "thread", "pool-d", "persist-d", "thread-q", "hybrid_pool-q"])
def gpu_ctx(request):
    """One device context per traversal-kernel variant: every GPU parity test runs against both."""
    import minipath_b200 as mp

    if mp.device_count() == 0:
        pytest.fail("no CUDA device: -m gpu tests need a B200 (there is no CPU fallback)")
    ctx = mp.Context(0)
    name, _, forced = request.param.partition("-")
    ctx.set_trace_variant({"octet": mp.TRACE_OCTET, "thread": mp.TRACE_THREAD, "persist": mp.TRACE_PERSIST, "hybrid": mp.TRACE_HYBRID, "pool": mp.TRACE_POOL, "hybrid_pool": mp.TRACE_HYBRID_POOL, "hybrid_pool_s0": mp.TRACE_HYBRID_POOL_S0}[name])
    # default node layout is AUTO (decoded f32 boxes / edges for thread-per-ray kernels, quantised records for persist / pool);
    # "-d" / "-q" force one layout for every kernel of the variant
    ctx.set_node_layout({"": mp.LAYOUT_AUTO, "d": mp.LAYOUT_DECODED, "q": mp.LAYOUT_QUANTISED}[forced])
    yield ctx
    ctx.close()
