"""Scenario compiler: slice_req dictionaries -> dense scenario-bank rows.

The reference keeps a scenario as a dict-of-dicts `slice_req` plus three association
matrices (`associations/mult_slice.py:58-347, 350-442`).  The device path wants plain
structs: one `irs_scenario` row (include/irs_b200.h) per network scenario, shared by every
env that is currently playing it.  This module is host-side only (NumPy).
"""
from __future__ import annotations

import numpy as np

MAX_UES = 25  # env_config/mult_slice.yml:14
MAX_SLICES = 5  # env_config/mult_slice.yml:13
NUM_RBS = 135  # env_config/mult_slice.yml:5
POISSON_K = 512

F_ACTIVE, F_REQ, F_THR, F_REL, F_LAT, F_EMBB, F_URLLC = 1, 2, 4, 8, 16, 32, 64

# agents/sched_colran.py:352-363: use case(s) of every slice type (SchedColORAN's reward)
USE_CASES = {
    "control_case_2": F_URLLC, "monitoring_case_1": F_EMBB, "robotic_surgery_case_1": F_EMBB | F_URLLC,
    "robotic_diagnosis": F_URLLC, "medical_monitoring": F_EMBB, "uav_app_case_1": F_EMBB,
    "uav_control_non_vlos": F_EMBB, "vr_gaming": F_EMBB | F_URLLC, "cloud_gaming": F_EMBB,
    "video_streaming_4k": F_EMBB,
}

# Restated from associations/mult_slice.py:58-347.  Columns:
# name, priority, reliability %, latency ms, throughput Mbps (None = intent not declared),
# buffer_size pkts, buffer_latency ms, message_size bits, traffic Mbps, min/max UEs.
SLICE_TYPES = [
    ("control_case_2", 1, 99.999999, 50, None, 10240, 100, 8192, 5, 4, 5),
    ("monitoring_case_1", 0, None, None, 10, 10240, 100, 8192, 10, 4, 5),
    ("robotic_surgery_case_1", 1, 99.9999, 20, 30, 1024000, 40, 16000, 30, 4, 5),
    ("robotic_diagnosis", 0, 99.999, 20, 15, 1024000, 40, 640, 15, 4, 5),
    ("medical_monitoring", 0, 99.9999, 100, 10, 10240, 200, 8000, 10, 4, 5),
    ("uav_app_case_1", 1, None, 200, 100, 1024000, 400, 65536, 100, 2, 4),
    ("uav_control_non_vlos", 1, 99.99, 140, 20, 10240, 300, 65536, 20, 4, 5),
    ("vr_gaming", 0, 99.99, 10, 100, 1024000, 20, 65536, 100, 2, 4),
    ("cloud_gaming", 0, None, 80, 50, 10240, 160, 65536, 50, 2, 5),
    ("video_streaming_4k", 0, None, None, 30, 10240, 100, 65536, 30, 2, 5),
]
SLICE_TYPE_NAMES = [t[0] for t in SLICE_TYPES]
# "mobility" (km/h) of the UEs of each type: carried in the slice_req dicts (history files, agent callbacks) for
# fidelity with the reference's files; nothing on the path reads it (mobilities/simple.py is constant)
SLICE_MOBILITY = [0, 72, 0, 0, 0, 30, 30, 0, 0, 0]

SLICE_DTYPE = np.dtype(
    [
        ("priority", "<f8"), ("traffic", "<f8"), ("buffer_size", "<f8"), ("buffer_latency", "<f8"),
        ("message_size", "<f8"), ("thr_value", "<f8"), ("rel_value", "<f8"), ("lat_value", "<f8"),
        ("flags", "<i4"), ("n_ues", "<i4"), ("lam_idx", "<i4"), ("pad", "<i4"),
    ]
)
SCENARIO_DTYPE = np.dtype([("sl", SLICE_DTYPE, (MAX_SLICES,)), ("ue_slice", "i1", (32,))])
assert SLICE_DTYPE.itemsize == 80 and SCENARIO_DTYPE.itemsize == 432


def slice_req_from_type(type_id: int) -> dict:
    """The reference-shaped slice_req entry of one slice type (operators as strings)."""
    name, prio, rel, lat, thr, bsz, blat, msg, traffic, umin, umax = SLICE_TYPES[type_id]
    params = {}
    if rel is not None:
        params[f"par{len(params)+1}"] = {"name": "reliability", "value": rel, "unit": "rate", "operator": "at_least"}
    if lat is not None:
        params[f"par{len(params)+1}"] = {"name": "latency", "value": lat, "unit": "ms", "operator": "at_most"}
    if thr is not None:
        params[f"par{len(params)+1}"] = {"name": "throughput", "value": thr, "unit": "Mbps", "operator": "at_least"}
    return {
        "name": name, "priority": prio, "parameters": params,
        "ues": {"buffer_size": bsz, "buffer_latency": blat, "message_size": msg, "mobility": SLICE_MOBILITY[type_id],
                "traffic": traffic, "min_number_ues": umin, "max_number_ues": umax},
    }


def _operator_name(op) -> str:
    if isinstance(op, str):
        return op
    return {"greater_equal": "at_least", "less_equal": "at_most"}.get(getattr(op, "__name__", ""), "unsupported")


def _philox4x32_7(counter, k0, k1):
    """Philox4x32-7 on Python ints (csrc/irs_rng.cuh philox4x32_7): host mirror for the few places that must predict
    a device draw (the shuffled episode order below)."""
    c = [int(x) & 0xFFFFFFFF for x in counter]
    k0, k1 = int(k0) & 0xFFFFFFFF, int(k1) & 0xFFFFFFFF
    for _ in range(7):
        p0, p1 = 0xD2511F53 * c[0], 0xCD9E8D57 * c[2]
        c = [((p1 >> 32) ^ c[1] ^ k0) & 0xFFFFFFFF, p1 & 0xFFFFFFFF, ((p0 >> 32) ^ c[3] ^ k1) & 0xFFFFFFFF, p0 & 0xFFFFFFFF]
        k0, k1 = (k0 + 0x9E3779B9) & 0xFFFFFFFF, (k1 + 0xBB67AE85) & 0xFFFFFFFF
    return c


def _mix32(x):
    x &= 0xFFFFFFFF
    x ^= x >> 16
    x = (x * 0x7FEB352D) & 0xFFFFFFFF
    x ^= x >> 15
    x = (x * 0x846CA68B) & 0xFFFFFFFF
    x ^= x >> 16
    return x


def shuffled_episode(seed: int, gid: int, gen: int, base: int, n: int, pos: int) -> int:
    """Element `pos` of the episode permutation an env plays with enable_random_episodes (comm_env.py:219-226,
    434-444) - the host mirror of csrc/irs_rng.cuh shuffled_episode (8-round Feistel network, cycle walking)."""
    K_PERM = 5
    k = _philox4x32_7([K_PERM << 16, gen, 0, gid], seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    if n <= 1:
        return base
    bits = 1
    while (1 << bits) < n:
        bits += 1
    bits += bits & 1
    half = bits >> 1
    mask = (1 << half) - 1
    x = pos
    while True:
        l, r = x >> half, x & mask
        for j in range(8):
            l, r = r, l ^ (_mix32(r ^ k[j & 3] ^ ((j * 0x9E3779B9) & 0xFFFFFFFF)) & mask)
        x = (l << half) | r
        if x < n:
            return base + x


_OPERATORS = {"at_least": np.greater_equal, "at_most": np.less_equal}


def callable_req(slice_req: dict) -> dict:
    """A slice_req dict with the operators as the reference holds them in memory (NumPy ufuncs,
    associations/mult_slice.py:48-55), for agent callbacks that call parameter["operator"](metric, value)
    (agents/common.py:124-135).  Deep copy; names given as strings are mapped."""
    out = {}
    for k, v in slice_req.items():
        if not v:
            out[k] = {}
            continue
        w = dict(v)
        w["ues"] = dict(v["ues"])
        w["parameters"] = {pk: dict(pv, operator=_OPERATORS.get(pv["operator"], pv["operator"]))
                           for pk, pv in v.get("parameters", {}).items()}
        out[k] = w
    return out


def req_from_row(row) -> tuple[dict, np.ndarray]:
    """(slice_req, slice_ue_assoc [5, 25]) of a dense scenario row - what Association.step returns for it
    (association.py:60-109).  Used where the row was generated on the device (assoc_mode 2) and no dict exists."""
    sua = np.zeros((MAX_SLICES, MAX_UES))
    for u in range(MAX_UES):
        if row["ue_slice"][u] >= 0:
            sua[int(row["ue_slice"][u]), u] = 1
    req = {}
    for s in range(MAX_SLICES):
        sl = row["sl"][s]
        if not (int(sl["flags"]) & F_REQ):
            req[f"slice_{s}"] = {}
            continue
        params = {}
        for flag, name, key, unit, op in ((F_REL, "reliability", "rel_value", "rate", "at_least"),
                                         (F_LAT, "latency", "lat_value", "ms", "at_most"),
                                         (F_THR, "throughput", "thr_value", "Mbps", "at_least")):
            if int(sl["flags"]) & flag:
                params[f"par{len(params) + 1}"] = {"name": name, "value": float(sl[key]), "unit": unit, "operator": op}
        ues = {"buffer_size": int(sl["buffer_size"]), "buffer_latency": int(sl["buffer_latency"]),
               "message_size": int(sl["message_size"]), "traffic": float(sl["traffic"])}
        name = "custom"
        for t in SLICE_TYPES:  # the slice type with these parameters, when there is one
            ref = slice_req_from_type(SLICE_TYPE_NAMES.index(t[0]))
            if all(ref["ues"][k] == ues[k] for k in ues) and ref["priority"] == sl["priority"] and \
                    {p["name"]: p["value"] for p in ref["parameters"].values()} == {p["name"]: p["value"] for p in params.values()}:
                name = t[0]
                ues = dict(ref["ues"])
                break
        req[f"slice_{s}"] = {"name": name, "priority": int(sl["priority"]), "parameters": params, "ues": ues}
    return req, sua


class ScenarioBank:
    """Dense table of scenarios + the Poisson CDF tables their traffic classes need."""

    def __init__(self):
        self.rows = []
        self.lambdas: list[float] = []
        self.reqs: list[dict] = []          # the slice_req dicts as given (history files, metrics.py:43-60)
        self.assocs: list[np.ndarray] = []  # slice_ue_assoc [5, 25] of every row

    def _lam_idx(self, lam: float) -> int:
        lam = float(lam)
        if lam not in self.lambdas:
            self.lambdas.append(lam)
        return self.lambdas.index(lam)

    def add(self, slice_req: dict, slice_ue_assoc: np.ndarray, basestation_slice_assoc=None) -> int:
        """slice_req: {"slice_0": {...} or {}, ...} as returned by Association.step
        (`sixg_radio_mgmt/sixg_radio_mgmt/association.py:60-109`)."""
        row = np.zeros((), SCENARIO_DTYPE)
        row["ue_slice"][:] = -1
        sua = np.asarray(slice_ue_assoc)
        assert sua.shape == (MAX_SLICES, MAX_UES)
        if (sua.sum(axis=0) > 1).any():
            raise ValueError("a UE is associated with more than one slice")
        bsa = None if basestation_slice_assoc is None else np.asarray(basestation_slice_assoc).reshape(-1)
        for s in range(MAX_SLICES):
            req = slice_req.get(f"slice_{s}", {})
            sl = row["sl"][s]
            ues = np.flatnonzero(sua[s])
            if len(ues) > 5:
                raise ValueError("at most 5 UEs per slice (agents/ib_sched.py:50)")
            row["ue_slice"][ues] = s
            sl["n_ues"] = len(ues)
            flags = 0
            active = (req != {}) if bsa is None else bool(bsa[s])
            if active:
                flags |= F_ACTIVE
            if req != {}:
                flags |= F_REQ | USE_CASES.get(req.get("name"), F_EMBB)
                ue = req["ues"]
                sl["priority"] = req["priority"]
                sl["traffic"] = ue["traffic"]
                sl["buffer_size"] = ue["buffer_size"]
                sl["buffer_latency"] = ue["buffer_latency"]
                sl["message_size"] = ue["message_size"]
                sl["lam_idx"] = self._lam_idx(ue["traffic"])
                for par in req["parameters"].values():
                    op = _operator_name(par["operator"])
                    want = {"throughput": "at_least", "reliability": "at_least", "latency": "at_most"}[par["name"]]
                    if op != want:
                        raise ValueError(f"unsupported intent operator {op} for {par['name']}")
                    if par["name"] == "throughput":
                        flags |= F_THR
                        sl["thr_value"] = par["value"]
                    elif par["name"] == "reliability":
                        flags |= F_REL
                        sl["rel_value"] = par["value"]
                    else:
                        flags |= F_LAT
                        sl["lat_value"] = par["value"]
            elif len(ues):
                # UEs in a slice without requirements keep the env defaults (comm_env.py:184-198)
                sl["buffer_size"], sl["buffer_latency"], sl["message_size"] = 10240, 100, 65536
            sl["flags"] = flags
        self.rows.append(row)
        self.reqs.append(slice_req)
        self.assocs.append(sua.astype(np.float64))
        return len(self.rows) - 1

    def add_typed(self, slice_type, ue_slice) -> int:
        """slice_type[5]: type id per slot (-1 unused); ue_slice[25]: slot of each UE (-1 none)."""
        req = {f"slice_{s}": (slice_req_from_type(int(t)) if t >= 0 else {}) for s, t in enumerate(slice_type)}
        sua = np.zeros((MAX_SLICES, MAX_UES))
        for u, s in enumerate(ue_slice):
            if s >= 0:
                sua[int(s), u] = 1
        return self.add(req, sua)

    @classmethod
    def templates(cls) -> "ScenarioBank":
        """Bank for assoc_mode 2 (scenario generated on the device per env and episode): row t is the
        TEMPLATE of slice type t - sl[0] holds the type's parameters, sl[0].n_ues its max and sl[0].pad
        its min number of UEs (associations/mult_slice.py:58-347)."""
        bank = cls()
        for t in range(len(SLICE_TYPES)):
            bank.add({"slice_0": slice_req_from_type(t)}, np.zeros((MAX_SLICES, MAX_UES)))
            row = bank.rows[-1]
            row["sl"][0]["n_ues"] = SLICE_TYPES[t][10]
            row["sl"][0]["pad"] = SLICE_TYPES[t][9]
        return bank

    def max_depth(self) -> int:
        arr = self.array()
        return int(arr["sl"]["buffer_latency"].max()) + 1

    def array(self) -> np.ndarray:
        return np.stack(self.rows) if self.rows else np.zeros((0,), SCENARIO_DTYPE)

    def poisson_cdf(self) -> np.ndarray:
        """[n_lambda, POISSON_K] float64; sample = #{j : cdf[j] <= u}.  Built once on the host
        and handed to the device (and, in tests, to the oracle) as data."""
        out = np.ones((max(1, len(self.lambdas)), POISSON_K))
        for i, lam in enumerate(self.lambdas):
            if lam + 12 * np.sqrt(lam) + 10 > POISSON_K:
                raise ValueError(f"traffic lambda {lam} too large for the {POISSON_K}-entry table")
            p = np.exp(-lam)
            c = 0.0
            for j in range(POISSON_K):
                c += p
                out[i, j] = min(c, 1.0)
                p = p * lam / (j + 1)
        return out


def generate_scenarios(n: int, seed: int = 10) -> tuple[np.ndarray, np.ndarray]:
    """Native generator with the distribution of associations/mult_slice.py:359-416
    (3..5 slices, slots and types without replacement, UEs/slice uniform in [min,max],
    random UE ids).  Not stream-compatible with gen_assoc_mult_slice.py."""
    rng = np.random.default_rng(seed)
    st = -np.ones((n, MAX_SLICES), np.int8)
    us = -np.ones((n, MAX_UES), np.int8)
    for i in range(n):
        k = int(rng.integers(3, MAX_SLICES, endpoint=True))
        slots = rng.choice(MAX_SLICES, k, replace=False)
        types = rng.choice(len(SLICE_TYPES), k, replace=False)
        counts = [int(rng.integers(SLICE_TYPES[t][9], SLICE_TYPES[t][10], endpoint=True)) for t in types]
        ues = rng.choice(MAX_UES, int(np.sum(counts)), replace=False)
        off = 0
        for slot, t, c in zip(slots, types, counts):
            st[i, slot] = t
            us[i, ues[off:off + c]] = slot
            off += c
    return st, us


def decode_row(row) -> tuple[np.ndarray, np.ndarray]:
    """One bank row -> (slice type id per slot, -1 = unused; slice of each UE) by matching the row's slice
    parameters against the slice-type table."""
    st = -np.ones(MAX_SLICES, np.int8)
    for s in range(MAX_SLICES):
        sl = row["sl"][s]
        if not (int(sl["flags"]) & F_REQ):
            continue
        for t, ty in enumerate(SLICE_TYPES):
            if (sl["priority"], sl["buffer_size"], sl["buffer_latency"], sl["message_size"], sl["traffic"]) == \
                    (ty[1], ty[5], ty[6], ty[7], ty[8]) and \
                    (ty[2] is None) == (not int(sl["flags"]) & F_REL) and (ty[4] is None) == (not int(sl["flags"]) & F_THR):
                st[s] = t
                break
    return st, np.array(row["ue_slice"][:MAX_UES], np.int8)


def export_reference_episode(directory: str, episode: int, slice_type, ue_slice, number_steps: int = 1000) -> str:
    """Write `ep_<episode>.npz` in the format of gen_assoc_mult_slice.py:229-237, which
    MultSliceAssociation.load_episode_data (associations/mult_slice.py:490-508) replays: per-step copies of the
    association matrices and of the slice_req dict (operators as numpy ufuncs, like the reference)."""
    import os

    ops = {"at_least": np.greater_equal, "at_most": np.less_equal}
    req = {}
    for s, t in enumerate(slice_type):
        r = slice_req_from_type(int(t)) if t >= 0 else {}
        for par in (r.get("parameters", {}) if r else {}).values():
            par["operator"] = ops[par["operator"]]
        req[f"slice_{s}"] = r
    sua = np.zeros((MAX_SLICES, MAX_UES))
    for u, s in enumerate(ue_slice[:MAX_UES]):
        if s >= 0:
            sua[int(s), u] = 1
    bsa = np.array([[1.0 if slice_type[s] >= 0 else 0.0 for s in range(MAX_SLICES)]])
    bua = np.array([sua.sum(axis=0)])
    rep = lambda a: np.repeat(np.asarray(a)[None], number_steps, axis=0)
    hist_req = np.empty(number_steps, object)
    to_use = np.empty(number_steps, object)
    used = np.flatnonzero(np.asarray(slice_type) >= 0)
    for k in range(number_steps):
        hist_req[k] = req
        to_use[k] = used
    os.makedirs(directory, exist_ok=True)
    path = os.path.join(directory, f"ep_{episode}.npz")
    np.savez_compressed(path, hist_basestation_ue_assoc=rep(bua), hist_basestation_slice_assoc=rep(bsa),
                        hist_slice_ue_assoc=rep(sua), hist_slice_req=hist_req,
                        hist_slices_lifetime=np.zeros((number_steps, MAX_SLICES)), hist_slices_to_use=to_use)
    return path
