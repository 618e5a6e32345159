"""Point-cloud wire format of the HT model (second half): the binary PLY written by
`save_ply` / read by `load_ply` (scene/gaussian_model_ht.py:411-446, :476-530) -- the file every downstream 3DGS
tool reads -- and the `capture()` tuple order (:90-105).

Layout (one `vertex` element, all properties `float`, little endian, in this order):
    x y z | nx ny nz (zeros) | f_dc_0..2 | f_rest_0..44 | opacity | scale_0..2 | rot_0..3        = 62 floats
`f_dc_*` / `f_rest_*` are CHANNEL-major: the stored `_features_dc [P,1,3]` / `_features_rest [P,15,3]` are transposed
to [P,3,1] / [P,3,15] and flattened (`.transpose(1, 2).flatten(start_dim=1)`, :425-428); all values are the stored,
un-activated parameters.

The reference builds the rows with `list(map(tuple, attributes))` -- a Python loop over every Gaussian -- and needs
the `plyfile` package.  Here the [P,62] matrix is assembled by torch on whatever device the parameters live on (one
gather on the GPU, one device->host copy) and written with a plain header + `ndarray.tofile`; reading is a header
parse + `np.fromfile`.  No dependency on plyfile.
"""
import os

import numpy as np
import torch

CAPTURE_FIELDS = ("active_sh_degree", "_xyz", "_features_dc", "_features_rest", "_scaling", "_rotation", "_opacity",
                  "max_radii2D", "xyz_gradient_accum", "denom", "optimizer_state_dict", "spatial_lr_scale", "P")


def list_of_attributes(n_dc=3, n_rest=45, n_scale=3, n_rot=4):
    """construct_list_of_attributes (scene/gaussian_model_ht.py:411-423)."""
    names = ["x", "y", "z", "nx", "ny", "nz"]
    names += [f"f_dc_{i}" for i in range(n_dc)]
    names += [f"f_rest_{i}" for i in range(n_rest)]
    names.append("opacity")
    names += [f"scale_{i}" for i in range(n_scale)]
    names += [f"rot_{i}" for i in range(n_rot)]
    return names


def pack_rows(xyz, features_dc, features_rest, opacity, scaling, rotation):
    """[P, 6 + 3 + 3*(M-1) + 1 + 3 + 4] float32 matrix in file order, on the inputs' device."""
    P = xyz.shape[0]
    for t, n in ((features_dc, "features_dc"), (features_rest, "features_rest"), (opacity, "opacity"), (scaling, "scaling"),
                 (rotation, "rotation")):
        if t.shape[0] != P:
            raise ValueError(f"{n} has {t.shape[0]} rows, xyz has {P}")
    f_dc = features_dc.detach().transpose(1, 2).flatten(start_dim=1)
    f_rest = features_rest.detach().transpose(1, 2).flatten(start_dim=1)
    xyz = xyz.detach()
    return torch.cat((xyz, torch.zeros_like(xyz), f_dc, f_rest, opacity.detach().reshape(P, 1), scaling.detach(),
                      rotation.detach()), dim=1).to(torch.float32).contiguous()


def save_ply(path, xyz, features_dc, features_rest, opacity, scaling, rotation):
    """Same bytes as `GaussianModel.save_ply` produces through plyfile (binary_little_endian 1.0, float properties)."""
    rows = pack_rows(xyz, features_dc, features_rest, opacity, scaling, rotation)
    names = list_of_attributes(features_dc.shape[1] * features_dc.shape[2], features_rest.shape[1] * features_rest.shape[2],
                               scaling.shape[1], rotation.shape[1])
    assert rows.shape[1] == len(names)
    d = os.path.dirname(path)
    if d:
        os.makedirs(d, exist_ok=True)
    header = "ply\nformat binary_little_endian 1.0\n" + f"element vertex {rows.shape[0]}\n" + \
             "".join(f"property float {n}\n" for n in names) + "end_header\n"
    with open(path, "wb") as f:
        f.write(header.encode("ascii"))
        rows.cpu().numpy().astype("<f4", copy=False).tofile(f)


def read_ply(path):
    """{property name: float32 [P]} of the first element of a binary little-endian PLY with float properties."""
    with open(path, "rb") as f:
        if f.readline().strip() != b"ply":
            raise ValueError(f"{path}: not a PLY file")
        fmt, count, names, in_vertex = None, None, [], False
        while True:
            line = f.readline()
            if not line:
                raise ValueError(f"{path}: truncated header")
            tok = line.decode("ascii").split()
            if not tok or tok[0] == "comment":
                continue
            if tok[0] == "format":
                fmt = tok[1]
            elif tok[0] == "element":
                if count is None:
                    count, in_vertex = int(tok[2]), True
                else:
                    in_vertex = False
            elif tok[0] == "property" and in_vertex:
                if tok[1] not in ("float", "float32"):
                    raise ValueError(f"{path}: property {tok[-1]} has type {tok[1]}, expected float")
                names.append(tok[2])
            elif tok[0] == "end_header":
                break
        if fmt != "binary_little_endian":
            raise ValueError(f"{path}: format {fmt} not supported (the reference writes binary_little_endian)")
        data = np.fromfile(f, dtype="<f4", count=count * len(names))
    if data.size != count * len(names):
        raise ValueError(f"{path}: expected {count} x {len(names)} floats, file holds {data.size}")
    data = data.reshape(count, len(names))
    return {n: data[:, i] for i, n in enumerate(names)}


def load_ply(path, max_sh_degree=3, num_gauss=-1, device="cpu"):
    """`GaussianModel.load_ply` (scene/gaussian_model_ht.py:476-530): returns the six stored parameter tensors
    {xyz [P,3], features_dc [P,1,3], features_rest [P,(D+1)^2-1,3], opacity [P,1], scaling [P,3], rotation [P,4]}."""
    cols = read_ply(path)
    n = len(cols["x"]) if num_gauss == -1 else num_gauss

    def numbered(prefix):
        ks = sorted((k for k in cols if k.startswith(prefix)), key=lambda k: int(k.split("_")[-1]))
        return np.stack([cols[k] for k in ks], axis=1) if ks else np.zeros((len(cols["x"]), 0), np.float32)

    rest = numbered("f_rest_")
    if rest.shape[1] != 3 * (max_sh_degree + 1) ** 2 - 3:
        raise ValueError(f"{path}: {rest.shape[1]} f_rest properties, expected {3 * (max_sh_degree + 1) ** 2 - 3}")
    out = {
        "xyz": np.stack((cols["x"], cols["y"], cols["z"]), axis=1),
        "features_dc": np.stack((cols["f_dc_0"], cols["f_dc_1"], cols["f_dc_2"]), axis=1)[:, :, None].transpose(0, 2, 1),
        "features_rest": rest.reshape(rest.shape[0], 3, rest.shape[1] // 3).transpose(0, 2, 1),
        "opacity": cols["opacity"][:, None],
        "scaling": numbered("scale_"),
        "rotation": numbered("rot"),
    }
    return {k: torch.tensor(np.ascontiguousarray(v[:n]), dtype=torch.float32, device=device) for k, v in out.items()}
