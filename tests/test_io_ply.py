"""PLY wire format (htgs_io) against the reference's save_ply / load_ply definitions
(scene/gaussian_model_ht.py:411-446, :476-530): attribute order, channel-major SH flattening, round trip."""
import numpy as np
import pytest
import torch

import htgs_io


def _params(P, seed=0):
    g = torch.Generator().manual_seed(seed)
    return dict(xyz=torch.randn(P, 3, generator=g), features_dc=torch.randn(P, 1, 3, generator=g),
                features_rest=torch.randn(P, 15, 3, generator=g), opacity=torch.randn(P, 1, generator=g),
                scaling=torch.randn(P, 3, generator=g), rotation=torch.randn(P, 4, generator=g))


def test_attribute_order_matches_reference_list():
    names = htgs_io.list_of_attributes()
    assert len(names) == 62
    assert names[:9] == ["x", "y", "z", "nx", "ny", "nz", "f_dc_0", "f_dc_1", "f_dc_2"]
    assert names[9] == "f_rest_0" and names[53] == "f_rest_44"
    assert names[54:] == ["opacity", "scale_0", "scale_1", "scale_2", "rot_0", "rot_1", "rot_2", "rot_3"]


def test_file_bytes_follow_the_reference_expressions(tmp_path):
    p = _params(37)
    path = str(tmp_path / "sub" / "point_cloud.ply")
    htgs_io.save_ply(path, **p)
    raw = open(path, "rb").read()
    head, body = raw.split(b"end_header\n", 1)
    lines = head.decode().splitlines()
    assert lines[:3] == ["ply", "format binary_little_endian 1.0", "element vertex 37"]
    assert lines[3:] == [f"property float {n}" for n in htgs_io.list_of_attributes()]
    got = np.frombuffer(body, dtype="<f4").reshape(37, 62)
    # the reference's own expressions (save_ply :425-441)
    f_dc = p["features_dc"].transpose(1, 2).flatten(start_dim=1).contiguous().numpy()
    f_rest = p["features_rest"].transpose(1, 2).flatten(start_dim=1).contiguous().numpy()
    xyz = p["xyz"].numpy()
    exp = np.concatenate((xyz, np.zeros_like(xyz), f_dc, f_rest, p["opacity"].numpy(), p["scaling"].numpy(), p["rotation"].numpy()), axis=1)
    assert np.array_equal(got, exp)


@pytest.mark.parametrize("P", [0, 1, 1000])
def test_round_trip(tmp_path, P):
    p = _params(P, seed=P)
    path = str(tmp_path / "pc.ply")
    htgs_io.save_ply(path, **p)
    q = htgs_io.load_ply(path, max_sh_degree=3)
    for k in p:
        assert q[k].shape == p[k].shape and torch.equal(q[k], p[k]), k
    half = htgs_io.load_ply(path, num_gauss=P // 2)
    assert half["xyz"].shape[0] == P // 2


def test_loader_rejects_wrong_degree_and_ascii(tmp_path):
    p = _params(5)
    path = str(tmp_path / "pc.ply")
    htgs_io.save_ply(path, **p)
    with pytest.raises(ValueError):
        htgs_io.load_ply(path, max_sh_degree=2)
    bad = str(tmp_path / "ascii.ply")
    open(bad, "w").write("ply\nformat ascii 1.0\nelement vertex 0\nproperty float x\nend_header\n")
    with pytest.raises(ValueError):
        htgs_io.read_ply(bad)


def test_capture_field_order():
    assert htgs_io.CAPTURE_FIELDS[:7] == ("active_sh_degree", "_xyz", "_features_dc", "_features_rest", "_scaling", "_rotation", "_opacity")
    assert len(htgs_io.CAPTURE_FIELDS) == 13
