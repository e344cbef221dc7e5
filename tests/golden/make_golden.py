#!/usr/bin/env python
"""Regenerates tests/golden/* from the mounted reference tree (/root/reference) and the CPU oracle.

Run in the build container only (the GPU box has no /root/reference); the outputs are committed.

  meshes.npz           ground / ben as welded (vertices, triangles) read from the reference's binary STL files
                       (mesh/ground.stl, mesh/ben.stl).  Welding = exact-duplicate merge in first-occurrence
                       order, which is what pcl::io::loadPolygonFileSTL (vtkSTLReader, merging on) yields:
                       98 / 162 and 2823 / 5489 (SRC/test/EmbreeTracer_test.cpp:89-90).
  ground.stl           the reference's 8 KB binary STL re-emitted triangle by triangle (loader fixture).
  ben_head.obj         first faces of ben.obj re-emitted in the Meshlab "vn/v ... f a//a" dialect (loader fixture).
  xt32_lidar_000{0,1}.json, transform_lidar_000{0,1}.json, trajectory.json
                       the reference's config VALUES re-serialised (own formatting, // comments kept in the
                       grammar because the reference's files have them: config/hesai-pandar-XT-32-lidar_0000.json:78).
  golden_xt32.npz      oracle outputs: per-ray t / triangle id for lidar_0000 and lidar_0001 on ground and
                       ground+ben (sensor-frame mode), point counts (1668 / 1781 reference-pinned,
                       SRC/test/EmbreeTracer_test.cpp:134, SRC/test/OptixTracer_test.cpp:168).
"""
import json
import os
import re
import struct
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
sys.path.insert(0, ROOT)


def read_stl(path):
    raw = open(path, "rb").read()
    (n,) = struct.unpack_from("<I", raw, 80)
    rec = np.frombuffer(raw, dtype=np.dtype([("n", "<f4", 3), ("v", "<f4", (3, 3)), ("a", "<u2")]), count=n, offset=84)
    return raw[:80], rec["n"].copy(), rec["v"].copy()


def weld(tri_verts):
    flat = tri_verts.reshape(-1, 3)
    seen, verts, idx = {}, [], np.empty(flat.shape[0], np.uint32)
    for i, p in enumerate(flat):
        k = p.tobytes()
        j = seen.get(k)
        if j is None:
            j = seen[k] = len(verts)
            verts.append(p)
        idx[i] = j
    return np.asarray(verts, np.float32), idx.reshape(-1, 3)


def strip_comments(text):
    out, i, in_str = [], 0, False
    while i < len(text):
        c = text[i]
        if in_str:
            out.append(c)
            if c == "\\":
                out.append(text[i + 1]); i += 1
            elif c == '"':
                in_str = False
        elif c == '"':
            in_str = True; out.append(c)
        elif text.startswith("//", i):
            while i < len(text) and text[i] != "\n":
                i += 1
            continue
        else:
            out.append(c)
        i += 1
    return "".join(out)


def load_json(path):
    return json.loads(strip_comments(open(path).read()))


def dump_with_comment(obj, path, comment):
    body = json.dumps(obj, indent=1)
    # one // comment inside the object and one trailing, so the product parser's comment handling is exercised
    body = body.replace("{\n", "{\n // " + comment + "\n", 1)
    open(path, "w").write(body + "\n// end of fixture\n")


def main():
    from oracle import lso

    meshes = {}
    for name in ("ground", "ben"):
        _, _, tv = read_stl(f"{REF}/mesh/{name}.stl")
        v, t = weld(tv)
        meshes[name + "_v"], meshes[name + "_t"] = v, t
        print(name, v.shape, t.shape)
    assert meshes["ground_v"].shape == (98, 3) and meshes["ground_t"].shape == (162, 3)
    assert meshes["ben_v"].shape == (2823, 3) and meshes["ben_t"].shape == (5489, 3)
    np.savez_compressed(f"{HERE}/meshes.npz", **meshes)

    # loader fixtures (re-emitted, not byte copies)
    hdr, nrm, tv = read_stl(f"{REF}/mesh/ground.stl")
    with open(f"{HERE}/ground.stl", "wb") as f:
        f.write(b"lidarshooter_b200 loader fixture (ground, 162 facets)".ljust(80, b" "))
        f.write(struct.pack("<I", tv.shape[0]))
        for n, tri in zip(nrm, tv):
            f.write(n.astype("<f4").tobytes() + tri.astype("<f4").tobytes() + b"\0\0")
    lines = [l for l in open(f"{REF}/mesh/ben.obj") if l[:2] in ("v ", "vn", "f ")]
    faces = [l for l in lines if l.startswith("f ")][:64]
    need = max(int(tok.split("/")[0]) for l in faces for tok in l.split()[1:])
    vs = [l for l in lines if l.startswith("v ")][:need]
    vns = [l for l in lines if l.startswith("vn")][:need]
    with open(f"{HERE}/ben_head.obj", "w") as f:
        f.write("# loader fixture: first 64 faces of ben (Meshlab dialect: vn/v interleaved, f a//a b//b c//c)\n")
        for a, b in zip(vns, vs):
            f.write(a); f.write(b)
        f.write("# %d vertices\n" % need)
        for l in faces:
            f.write(l)

    # configs
    cfgs = {}
    for uid in ("0000", "0001"):
        c = load_json(f"{REF}/config/hesai-pandar-XT-32-lidar_{uid}.json")
        c["device"]["sensorConfigFile"] = f"transform_lidar_{uid}.json"
        c["outputFolder"] = "."
        dump_with_comment(c, f"{HERE}/xt32_lidar_{uid}.json", "Units of channels.vertical: degrees above horizontal")
        t = load_json(f"{REF}/config/transform-lidar_{uid}.json")
        dump_with_comment(t, f"{HERE}/transform_lidar_{uid}.json", "sensor extrinsics")
        cfgs[uid] = c
    dump_with_comment(load_json(f"{REF}/config/trajectory.json"), f"{HERE}/trajectory.json", "piecewise-constant twists")

    # oracle golden vectors
    out = {}
    for uid, c in cfgs.items():
        sc = c["device"]["sensorConfig"]
        s = lso.Sensor(c["channels"]["vertical"], c["channels"]["horizontal"]["range"]["begin"],
                       c["channels"]["horizontal"]["range"]["end"], c["channels"]["horizontal"]["count"],
                       [sc["sensor_to_base"][k] for k in ("qw", "qx", "qy", "qz")],
                       [sc["base_to_origin"]["tx"], sc["base_to_origin"]["ty"], sc["sensor_to_base"]["tz"]])
        g = (meshes["ground_v"], meshes["ground_t"], lso.IDENTITY_AFFINE)
        b = (meshes["ben_v"], meshes["ben_t"], lso.IDENTITY_AFFINE)
        for tag, ms in (("ground", [g]), ("ground_ben", [g, b])):
            r = lso.scan(s, ms, mode=0, use_bvh=False)
            r2 = lso.scan(s, ms, mode=0, use_bvh=True)
            assert np.array_equal(r["tri"], r2["tri"]) and np.array_equal(r["t"], r2["t"]), "brute != bvh"
            out[f"l{uid}_{tag}_t"], out[f"l{uid}_{tag}_tri"] = r["t"], r["tri"]
            print(uid, tag, "points:", r["points"].shape[0])
        out[f"l{uid}_dirs"] = lso.ray_dirs(s.tables)
        out[f"l{uid}_Rinv"], out[f"l{uid}_R"] = s.Rinv, s.R
    assert (out["l0000_ground_tri"] >= 0).sum() == 1668     # reference-pinned
    assert (out["l0000_ground_ben_tri"] >= 0).sum() == 1781  # reference-pinned
    np.savez_compressed(f"{HERE}/golden_xt32.npz", **out)


if __name__ == "__main__":
    main()
