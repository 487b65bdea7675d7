"""Checkpoint files with the reference's names (``train.py:145,210-217``; ``sample.py:41-44``).

``save_dir/model.ckpt-<step>.npz`` holds what ``tf.train.Saver(tf.global_variables())`` saves:
every trainable variable under its TF name, Adam's ``m``/``v`` slots (``<name>/Adam``,
``<name>/Adam_1``), ``global_step`` and the ``lr`` variable; ``model.ckpt-<step>.index`` is a
JSON listing (name -> shape) so the reference CI's ``test -s model.ckpt-N.index`` still holds;
``checkpoint`` is the same text state file TF writes.  The TF binary tensor format itself is out
of scope (no reader or writer exists here) - DESIGN.md.
"""
import json
import os
import re

import numpy as np


def save(save_dir, state, step):
    prefix = os.path.join(save_dir, "model.ckpt-%d" % step)
    arrays = {}
    for name, v in state["variables"].items():
        arrays[name] = v
    for name, v in state.get("adam_m", {}).items():
        arrays[name + "/Adam"] = v
    for name, v in state.get("adam_v", {}).items():
        arrays[name + "/Adam_1"] = v
    arrays["global_step"] = np.asarray(state["global_step"], dtype=np.int64)
    arrays["lr"] = np.asarray(state["lr"], dtype=np.float32)
    tmp = prefix + ".tmp.npz"
    np.savez(tmp, **arrays)
    os.replace(tmp, prefix + ".npz")
    with open(prefix + ".index", "w") as f:
        json.dump({k: list(np.shape(v)) for k, v in arrays.items()}, f)
    paths = [p for p in _all_paths(save_dir) if p != os.path.basename(prefix)] + [os.path.basename(prefix)]
    with open(os.path.join(save_dir, "checkpoint"), "w") as f:
        f.write('model_checkpoint_path: "%s"\n' % os.path.basename(prefix))
        for p in paths:
            f.write('all_model_checkpoint_paths: "%s"\n' % p)
    return os.path.join(save_dir, "model.ckpt")


def _all_paths(save_dir):
    state = os.path.join(save_dir, "checkpoint")
    if not os.path.isfile(state):
        return []
    out = []
    for line in open(state):
        m = re.match(r'all_model_checkpoint_paths: "(.*)"', line.strip())
        if m and os.path.isfile(os.path.join(save_dir, os.path.basename(m.group(1)) + ".npz")):
            out.append(os.path.basename(m.group(1)))
    return out


def latest_checkpoint(save_dir):
    """``tf.train.latest_checkpoint``: prefix of the newest checkpoint, or None."""
    state = os.path.join(save_dir, "checkpoint")
    if not os.path.isfile(state):
        return None
    for line in open(state):
        m = re.match(r'model_checkpoint_path: "(.*)"', line.strip())
        if m:
            prefix = os.path.join(save_dir, os.path.basename(m.group(1)))
            return prefix if os.path.isfile(prefix + ".npz") else None
    return None


def load(prefix):
    z = np.load(prefix + ".npz")
    state = {"variables": {}, "adam_m": {}, "adam_v": {}}
    for k in z.files:
        if k == "global_step":
            state["global_step"] = int(z[k])
        elif k == "lr":
            state["lr"] = float(z[k])
        elif k.endswith("/Adam_1"):
            state["adam_v"][k[:-len("/Adam_1")]] = z[k]
        elif k.endswith("/Adam"):
            state["adam_m"][k[:-len("/Adam")]] = z[k]
        else:
            state["variables"][k] = z[k]
    if not state["adam_m"]:
        del state["adam_m"], state["adam_v"]
    return state
