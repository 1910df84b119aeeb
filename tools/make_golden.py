"""Generate the golden fixtures under tests/golden/ from the parts of the reference that ARE executable /
readable in this container (SURVEY.md §8c "weak pins"):

  quatmath.npz   euler2quat / quat2euler outputs of the reference's own utils/quatmath.py (numpy-only, imported
                 read-only from /root/reference) on seeded random inputs.
  dapg_<task>.npz  the four pretrained DAPG policies (algos/dapg_pretrained/*.pickle), stub-unpickled: layer
                 weights, input/output normalisation and the (obs_dim, act_dim) they pin.

Usage: python tools/make_golden.py [/root/reference]
"""
import importlib.util
import io
import os
import pickle
import sys

import numpy as np

ref = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")
os.makedirs(out, exist_ok=True)

# ---- quatmath
spec = importlib.util.spec_from_file_location("ref_quatmath", os.path.join(ref, "mj_envs_vision", "utils", "quatmath.py"))
qm = importlib.util.module_from_spec(spec)
spec.loader.exec_module(qm)
rng = np.random.default_rng(20240)
eul = rng.uniform(-np.pi, np.pi, (256, 3))
eul[:8, 1] = np.array([np.pi / 2, -np.pi / 2, 0, 0, 1e-9, np.pi / 2 - 1e-9, 0.3, -0.3])   # gimbal-lock edge cases
quat_in = rng.normal(size=(256, 4))
quat_in /= np.linalg.norm(quat_in, axis=1, keepdims=True)
quat_in[:4] *= np.array([[0.5], [2.0], [1.0], [1e-3]])                                        # un-normalised inputs
np.savez(os.path.join(out, "quatmath.npz"), euler_in=eul, euler2quat=qm.euler2quat(eul), quat_in=quat_in,
         quat2euler=qm.quat2euler(quat_in), pen_reset_e=np.stack([eul[:, 0] / np.pi, eul[:, 1] / np.pi, 0 * eul[:, 0]], 1),
         pen_reset_q=qm.euler2quat(np.stack([eul[:, 0] / np.pi, eul[:, 1] / np.pi, 0 * eul[:, 0]], 1)))


# ---- DAPG policies
class _Stub:
    def __init__(self, *a, **k):
        pass

    def __setstate__(self, st):
        self.__dict__.update(st if isinstance(st, dict) else {"state": st})


def _rebuild_tensor_v2(storage, offset, size, stride, *a):
    arr = np.asarray(storage)
    return np.lib.stride_tricks.as_strided(arr[offset:], shape=size, strides=[s * arr.itemsize for s in stride]).copy()


def _rebuild_parameter(data, requires_grad, hooks):
    return data


class U(pickle.Unpickler):
    def find_class(self, module, name):
        if module.startswith("mjrl") or module.startswith("torch.nn") or module == "collections" and name == "OrderedDict":
            if name == "OrderedDict":
                import collections
                return collections.OrderedDict
            return type(name, (_Stub,), {})
        if module == "torch._utils" and name == "_rebuild_tensor_v2":
            return _rebuild_tensor_v2
        if module == "torch._utils" and name == "_rebuild_parameter":
            return _rebuild_parameter
        if module == "torch" and name.endswith("Storage"):
            return name
        if module.startswith("numpy.core"):
            module = module.replace("numpy.core", "numpy._core")
        return super().find_class(module, name)

    def persistent_load(self, pid):
        raise pickle.UnpicklingError("unexpected persistent id")


def load_policy(path):
    with open(path, "rb") as f:
        data = f.read()
    try:
        return U(io.BytesIO(data)).load()
    except Exception:
        import torch  # torch-pickled tensors inside a plain pickle: let torch rebuild storages
        class TU(U):
            def find_class(self, module, name):
                if module.startswith("torch") and not module.startswith("torch.nn"):
                    return pickle.Unpickler.find_class(self, module, name)
                return super().find_class(module, name)
        return TU(io.BytesIO(data)).load()


def to_np(x):
    if hasattr(x, "detach"):
        return x.detach().cpu().numpy()
    return np.asarray(x)


for task in ("door", "hammer", "pen", "relocate"):
    pol = load_policy(os.path.join(ref, "mj_envs_vision", "algos", "dapg_pretrained", f"{task}-v0.pickle"))
    d = pol.__dict__
    model = d["model"].__dict__
    mods = model.get("_modules", {})
    layers = mods["fc_layers"].__dict__["_modules"]
    W, B = [], []
    for k in sorted(layers, key=int):
        p = layers[k].__dict__["_parameters"]
        W.append(to_np(p["weight"]).astype(np.float32))
        B.append(to_np(p["bias"]).astype(np.float32))
    g = lambda k, dflt: to_np(model[k]).astype(np.float32).reshape(-1) if k in model and model[k] is not None else dflt
    n, m_ = int(d["n"]), int(d["m"])
    save = dict(n=n, m=m_, log_std=to_np(d["log_std"]).astype(np.float32).reshape(-1),
                in_shift=g("in_shift", np.zeros(n, np.float32)), in_scale=g("in_scale", np.ones(n, np.float32)),
                out_shift=g("out_shift", np.zeros(m_, np.float32)), out_scale=g("out_scale", np.ones(m_, np.float32)))
    for i, (w, b) in enumerate(zip(W, B)):
        save[f"W{i}"] = w
        save[f"b{i}"] = b
    np.savez_compressed(os.path.join(out, f"dapg_{task}.npz"), **save)
    print(task, "obs", n, "act", m_, "layers", [w.shape for w in W])
