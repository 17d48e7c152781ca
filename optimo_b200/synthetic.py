"""Synthetic scenes and candidate batches of BASELINE.json's configs 4 and 5.

Harness data only: this module builds scene-JSON dicts and descriptor matrices; it evaluates nothing.
Random numbers follow the rule fixed in SURVEY.md: std::mt19937_64(seed), u() = (rng() >> 11) * 2**-53.
"""
import copy
import math

import numpy as np

_NN, _MM = 312, 156
_MATRIX_A = np.uint64(0xB5026F5AA96619E9)
_UM = np.uint64(0xFFFFFFFF80000000)
_LM = np.uint64(0x7FFFFFFF)


class MT19937_64:
    """std::mt19937_64, vectorised over a batch of seeds (one independent generator per seed)."""

    def __init__(self, seeds):
        seeds = np.atleast_1d(np.asarray(seeds, dtype=np.uint64))
        mt = np.zeros((_NN, seeds.size), dtype=np.uint64)
        mt[0] = seeds
        with np.errstate(over="ignore"):
            for i in range(1, _NN):
                prev = mt[i - 1]
                mt[i] = np.uint64(6364136223846793005) * (prev ^ (prev >> np.uint64(62))) + np.uint64(i)
        self.mt = mt
        self.idx = _NN

    def _twist(self):
        mt = self.mt
        for i in range(_NN):
            x = (mt[i] & _UM) | (mt[(i + 1) % _NN] & _LM)
            mt[i] = mt[(i + _MM) % _NN] ^ (x >> np.uint64(1)) ^ np.where(x & np.uint64(1), _MATRIX_A, np.uint64(0))
        self.idx = 0

    def next_u64(self):
        if self.idx >= _NN:
            self._twist()
        x = self.mt[self.idx].copy()
        self.idx += 1
        x ^= (x >> np.uint64(29)) & np.uint64(0x5555555555555555)
        x ^= (x << np.uint64(17)) & np.uint64(0x71D67FFFEDA60000)
        x ^= (x << np.uint64(37)) & np.uint64(0xFFF7EEE000000000)
        x ^= x >> np.uint64(43)
        return x

    def uniform(self):
        return (self.next_u64() >> np.uint64(11)).astype(np.float64) * (2.0 ** -53)


def iter_joints(joint):
    """DFS pre-order over a scene-JSON joint tree (Object::GetJoints order)."""
    yield joint
    for c in joint.get("children", []):
        yield from iter_joints(c)


def count_control_points(scene):
    n = 0
    for j in iter_joints(scene["joints"]):
        for key in ("rot_x", "rot_y", "rot_z"):
            n += len(j[key]["curve"])
    for h in scene.get("ik_handlers", []):
        for key in ("pos_x", "pos_y", "pos_z"):
            n += len(h[key]["curve"])
    return n


def stick_240(stick_scene, repeats=6, period=40.0):
    """Config 4 scene: the stick-figure skeleton over frames 1..240, every original control point
    (t, v, theta, t_b, t_f) replicated at t + 40 r, r = 0..5, appended replica by replica."""
    scene = copy.deepcopy(stick_scene)
    scene["min_frame"], scene["max_frame"] = 1, int(period * repeats)
    for j in iter_joints(scene["joints"]):
        for key in ("rot_x", "rot_y", "rot_z"):
            base = j[key]["curve"]
            j[key]["curve"] = [[cp[0] + period * r, cp[1], cp[2], cp[3], cp[4]] for r in range(repeats) for cp in base]
    return scene


def perturbed_candidates(x0, B, seed_base=0xB200, first=0):
    """Candidates c = first .. first+B-1 of the Config-4 rule: generator mt19937_64(seed_base + c); for every
    control point in descriptor order three draws (theta, t_b, t_f): theta += 0.6 (u1 - 0.5),
    t_b = -(0.5 + u2), t_f = +(0.5 + u3); t and v untouched."""
    x0 = np.asarray(x0, dtype=np.float64)
    n_cp = x0.size // 5
    rng = MT19937_64(np.arange(first, first + B, dtype=np.uint64) + np.uint64(seed_base))
    X = np.tile(x0, (B, 1))
    for k in range(n_cp):
        u1, u2, u3 = rng.uniform(), rng.uniform(), rng.uniform()
        X[:, 5 * k + 2] = x0[5 * k + 2] + 0.6 * (u1 - 0.5)
        X[:, 5 * k + 3] = -(0.5 + u2)
        X[:, 5 * k + 4] = +(0.5 + u3)
    return X


def tree_scene(n_joints=64, n_frames=2000, key_every=10, seed=64):
    """Config 5 scene: root + (n_joints - 1) joints, parent(j) = (j-1)//2 (binary tree); offset_j = (0.1 + 0.4 u)
    along a unit direction (uniform in the ball by rejection, rejected when within 5 degrees of -y); every joint
    and axis keyed every `key_every` frames, t = 1, 11, ..., v = 0.5 sin(0.05 t + phi_var), theta = 0,
    t_b = -3, t_f = +3. Joints are emitted so that the JSON's DFS pre-order is the construction order."""
    rng = MT19937_64([seed])

    def u():
        return float(rng.uniform()[0])

    cos_lim = math.cos(math.radians(5.0))
    offsets = [[0.0, 0.9, 0.0]]
    for _ in range(1, n_joints):
        while True:
            d = [2.0 * u() - 1.0, 2.0 * u() - 1.0, 2.0 * u() - 1.0]
            n = math.sqrt(d[0] ** 2 + d[1] ** 2 + d[2] ** 2)
            if n > 1.0 or n < 1e-3:
                continue
            d = [c / n for c in d]
            if -d[1] > cos_lim:
                continue
            break
        length = 0.1 + 0.4 * u()
        offsets.append([length * c for c in d])
    times = list(range(1, n_frames + 2, key_every))
    nodes = []
    var_index = 0
    for j in range(n_joints):
        node = {"name": "root" if j == 0 else f"j{j}", "offset": offsets[j], "children": []}
        for key in ("rot_x", "rot_y", "rot_z"):
            if j == 0:
                node[key] = {"var": 0.0, "curve": []}
            else:
                phi = 0.7 * var_index
                node[key] = {"var": 0.0, "curve": [[float(t), 0.5 * math.sin(0.05 * t + phi), 0.0, -3.0, 3.0] for t in times]}
                var_index += 1
        nodes.append(node)
    for j in range(1, n_joints):
        nodes[(j - 1) // 2]["children"].append(nodes[j])
    return {"min_frame": 1, "max_frame": n_frames, "frames_per_second": 30, "joints": nodes[0], "ik_handlers": []}


def flops_per_sample_eval(n_joints, bezier_evals_per_sample, newton_iters_per_sample, with_positions=False, ik_iters_per_sample=0.0):
    """SURVEY.md §8(d) ALGORITHMIC FLOPs per sample-eval:
    F_se = 393 (J - 1) + sum_bezier (45 + 14 k) [+ 92 J] [+ 450 it_IK]."""
    f = 393.0 * (n_joints - 1) + 45.0 * bezier_evals_per_sample + 14.0 * newton_iters_per_sample
    if with_positions:
        f += 92.0 * n_joints
    f += 450.0 * ik_iters_per_sample
    return f
