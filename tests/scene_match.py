"""Cross-walk between the reference's BVH leaves (ref_driver `scene` dump) and the flat scene's instance table."""
import numpy as np
from scipy.spatial import cKDTree

KIND = {"sphere": 0, "rect": 1, "box": 2}
PLANE_FROM_REF = {0: 0, 1: 1, 2: 2}  # plane_enum XY, XZ, YZ (primitive.h:11-16)


def ref_leaf_signature(leaf):
    prim = leaf["prim"]
    medium = prim["type"] == "volume"
    density = prim["density"] if medium else 0.0
    if medium:
        prim = prim["boundary"]
    kind = KIND[prim["type"]]
    if kind == 0:
        params = list(prim["center"]) + [prim["radius"], 0.0, 0.0]
    elif kind == 1:
        params = [prim["x0"], prim["z0"], prim["x1"], prim["z1"], prim["y"], float(prim["plane"])]
    else:
        params = list(prim["p0"]) + list(prim["p1"])
    return np.array(list(leaf["transform"]) + params + [kind, float(medium), density], dtype=np.float64)


def flat_signature(inst):
    kind = inst["prim_kind"]
    params = [float(x) for x in inst["params"]]
    if kind == 0:
        params = params[:4] + [0.0, 0.0]
    elif kind == 1:
        params = params[:5] + [float(inst["plane"])]
    medium = bool(inst["flags"] & 1)
    return np.array(list(inst["fwd"].ravel()) + params + [kind, float(medium), inst["density"] if medium else 0.0],
                    dtype=np.float64)


def match_leaves(ref_scene, flat):
    """Returns (ref_to_flat int array, max signature distance).  Identical twins are interchangeable; they are paired
    greedily in order so the mapping stays a bijection."""
    ref_sig = np.array([ref_leaf_signature(l) for l in ref_scene["leaves"]])
    flat_sig = np.array([flat_signature(i) for i in flat["instances"]])
    assert ref_sig.shape == flat_sig.shape, (ref_sig.shape, flat_sig.shape)
    tree = cKDTree(flat_sig)
    k = min(8, len(flat_sig))
    dist, idx = tree.query(ref_sig, k=k)
    dist = dist.reshape(len(ref_sig), -1)
    idx = idx.reshape(len(ref_sig), -1)
    used = np.zeros(len(flat_sig), dtype=bool)
    out = np.full(len(ref_sig), -1, dtype=np.int64)
    worst = 0.0
    for r in np.argsort(dist[:, 0], kind="stable"):
        for d, j in zip(dist[r], idx[r]):
            if not used[j]:
                used[j] = True
                out[r] = j
                worst = max(worst, float(d))
                break
        else:
            free = np.nonzero(~used)[0]
            dd = np.linalg.norm(flat_sig[free] - ref_sig[r], axis=1)
            j = free[np.argmin(dd)]
            used[j] = True
            out[r] = j
            worst = max(worst, float(dd.min()))
    return out, worst


def remap_ref_hits(ref_hits, ref_to_flat):
    out = ref_hits.copy()
    ok = ref_hits["instance"] >= 0
    out["instance"][ok] = ref_to_flat[ref_hits["instance"][ok]]
    return out
