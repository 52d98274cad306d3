"""CPU: the C++ host mirror of the reference's interface (hostapi/): camera, projection, OBJ loading. The matrices
must be bit-identical to the ones the reference's own Camera / perspectiveMatrix / inverse4x4 produced
(tests/golden/frames.npz holds them for five cameras and three aspect ratios)."""
import os

import numpy as np
import pytest

import triangles_sdf_cpu_raytracing_b200 as rt
from tests import util

ORBIT_DX = float(np.float32(2.0 * np.pi / 64.0 / 0.01))
CAMS = {
    "default": dict(),
    "orbit5": dict(n_rotate=5, dx=ORBIT_DX, dy=0.0),
    "orbit23tilt": dict(n_rotate=23, dx=ORBIT_DX, dy=1.5),
    "plusx": dict(eye=(2.5, 0.0, 0.0)),
    "close": dict(eye=(0.3, 0.2, 1.2)),
}


def test_camera_and_projection_match_the_reference_bit_for_bit(golden):
    seen = set()
    for key in golden.cases:
        c = golden.case(key)
        pos, _view, vinv = rt.host_camera(**CAMS[c["cam"]])
        assert util.same_bits(pos, c["rayPos"]), key
        assert util.same_bits(vinv, c["viewInv"]), key
        _proj, pinv = rt.host_projection(c["W"], c["H"])
        assert util.same_bits(pinv, c["projInv"]), key
        seen.add(c["cam"])
    assert seen == set(CAMS)


def test_obj_loader_rules(tmp_path):
    # quads become fans, (v, vt, vn) triples dedup in order of first use, negative indices count from the end
    p = tmp_path / "m.obj"
    p.write_text("v 0 0 0\nv 1 0 0\nv 1 1 0\nv 0 1 0\nvn 0 0 1\nvt 0 0\n"
                 "f 1/1/1 2/1/1 3/1/1 4/1/1\nf -4//1 -3//1 -2//1\nf 1 2 3\n")
    v, i = rt.host_load_obj(str(p), scale=False)
    assert i.tolist() == [0, 1, 2, 0, 2, 3, 4, 5, 6, 7, 8, 9]
    assert v.shape == (10, 4) and (v[:, 3] == 1).all()
    assert np.array_equal(v[4, :3], [0, 0, 0]) and np.array_equal(v[9, :3], [1, 1, 0])
    with pytest.raises(rt.RtError):
        rt.host_load_obj(str(tmp_path / "missing.obj"))


@pytest.mark.parametrize("name", util.MESHES)
def test_obj_loading_and_normalisation_match_the_reference(golden, name):
    path = f"/root/reference/resources/{name}.obj"
    if not os.path.exists(path):
        pytest.skip("reference resources are not on this machine")
    v, i = rt.host_load_obj(path, scale=True)
    a = golden.scene_arrays(name)
    assert np.array_equal(i, a["indices"])
    assert util.same_bits(v, a["vpos4"])
