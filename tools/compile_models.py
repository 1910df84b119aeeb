"""Compile the reference's MJCF scenes into the binary model blobs shipped under
mj_envs_b200/compiled/ (the reference tree does not exist on the GPU box).

Usage:  python tools/compile_models.py [/root/reference]
Reads  <ref>/mj_envs_vision/hand_manipulation_suite/assets/DAPG_<task>.xml (read-only).
"""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from mj_envs_b200.mjcf import load_mjcf, Model  # noqa: E402

ref = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
assets = os.path.join(ref, "mj_envs_vision", "hand_manipulation_suite", "assets")
out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "mj_envs_b200", "compiled")
os.makedirs(out, exist_ok=True)
for task in ("hammer", "door", "pen", "relocate"):
    m = load_mjcf(os.path.join(assets, f"DAPG_{task}.xml"), task)
    blob = m.to_blob()
    m2 = Model.from_blob(blob)
    assert m2.nv == m.nv and m2.names == m.names
    with open(os.path.join(out, f"{task}.adroit"), "wb") as f:
        f.write(blob)
    s = m.A["sizes"]
    print(f"{task}: nv={m.nv} nbody={m.nbody} ngeom={m.ngeom} nsite={m.nsite} nu={m.nu} ntendon={m.ntendon} "
          f"npair={m.npair} (dynamic {s[12]}) npar={m.npar} obs={m.obs_dim} blob={len(blob)}B "
          f"meaninertia={m.A['meaninertia'][0]:.6g}")
