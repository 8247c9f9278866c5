"""Synthetic inputs of the shapes BASELINE.json names, counter-based so that every
shard, rerun and the CPU oracle see identical rows without materialising CSV at the big sizes:

    u(i, j) = (splitmix64(seed * 2**40 + i * 1024 + j) >> 11) * 2**-53      row i, draw j

call_hangup restates the reference's generator python/app/call_hangup.py:1-59 (with
python/lib/util.py:107-111 selectRandomFromList and python/lib/sampler.py:367-399
GaussianRejectSampler); its python-2 `random` stream is unseeded, so only the distribution -- not a
particular file -- can be reproduced.
"""
from __future__ import annotations

import numpy as np

_U64 = np.uint64
DEFAULT_SEED = 20260101


def splitmix64(x: np.ndarray) -> np.ndarray:
    x = x.astype(np.uint64, copy=True)
    with np.errstate(over="ignore"):
        x += _U64(0x9E3779B97F4A7C15)
        z = x
        z = (z ^ (z >> _U64(30))) * _U64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> _U64(27))) * _U64(0x94D049BB133111EB)
        z = z ^ (z >> _U64(31))
    return z


def uniform(seed: int, rows: np.ndarray, draw: int) -> np.ndarray:
    """u(i, draw) in [0, 1) for the given row indices."""
    with np.errstate(over="ignore"):
        key = _U64((seed << 40) & 0xFFFFFFFFFFFFFFFF) + rows.astype(np.uint64) * _U64(1024) + _U64(draw)
    return (splitmix64(key) >> _U64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)


# ---- call_hangup (configs C1 / C2) -----------------------------------------------------------------
CUST_TYPES = ["business", "residence"]
RES_ISSUES = ["internet", "cable", "billing", "other"]
BIZ_ISSUES = ["internet", "billing", "other"]
TOD = ["AM", "PM"]
AREA_CODES = [408, 607, 336, 267, 646, 760, 615, 980, 828, 385, 941, 305, 971, 510, 574, 620, 507, 540, 206, 262, 847,
              941, 470, 323, 630, 615, 346, 216, 920, 903, 423, 614, 440, 419, 832, 678, 608, 678, 571, 248, 321, 301,
              630, 719, 209, 770, 615, 971, 937, 703]
ISSUE_CARD = ["internet", "cable", "billing", "other"]   # resource/call_hangup.json cardinality order
HOLD_POINTS = np.arange(60, 600, 60)                     # split grid of attr 5: 60..540


def call_hangup_fields(n: int, seed: int = DEFAULT_SEED, start: int = 0):
    """Raw field arrays of rows [start, start+n): dict with cust (0/1), issue (index into ISSUE_CARD),
    tod (0/1), hold (int), hung (0 = 'T', 1 = 'F'), area (int)."""
    rows = np.arange(start, start + n, dtype=np.uint64)
    cust = (uniform(seed, rows, 0) * 2).astype(np.int64)                  # randint(0, 1)
    u_issue = uniform(seed, rows, 1)
    res_issue = (u_issue * 4).astype(np.int64)                            # index into RES_ISSUES
    biz_issue = np.array([0, 2, 3])[(u_issue * 3).astype(np.int64)]       # internet, billing, other
    issue = np.where(cust == 1, res_issue, biz_issue)
    area = np.array(AREA_CODES)[(uniform(seed, rows, 2) * len(AREA_CODES)).astype(np.int64)]
    tod = (uniform(seed, rows, 3) * 2).astype(np.int64)
    # GaussianRejectSampler(mean, sd): x ~ U(mean-3sd, mean+3sd), y ~ U(0, 1.05 fmax), accept y < f(x)
    mean = np.where(tod == 0, 500.0, 400.0)
    sd = np.where(tod == 0, 80.0, 60.0)
    hold = np.zeros(n, dtype=np.float64)
    pending = np.ones(n, dtype=bool)
    attempt = 0
    while pending.any():
        idx = np.nonzero(pending)[0]
        r = rows[idx]
        x = mean[idx] - 3 * sd[idx] + uniform(seed, r, 16 + 2 * attempt) * 6 * sd[idx]
        fmax = 1.0 / (np.sqrt(2.0 * 3.14) * sd[idx])
        y = uniform(seed, r, 17 + 2 * attempt) * 1.05 * fmax
        f = fmax * np.exp(-(x - mean[idx]) ** 2 / (2.0 * sd[idx] ** 2))
        ok = y < f
        hold[idx[ok]] = x[ok]
        pending[idx[ok]] = False
        attempt += 1
        if attempt > 400:
            raise RuntimeError("rejection sampler did not converge")
    hold_i = hold.astype(np.int64)
    thr = np.full(n, 180, dtype=np.int64)
    biz = cust == 0
    thr[biz & (issue == 0)] = 450
    thr[biz & (issue == 2)] = 300
    thr[~biz & (issue == 0)] = 350
    thr[~biz & (issue == 2)] = 250
    rnd = (uniform(seed, rows, 4) * 101).astype(np.int64)                 # randint(0, 100) inclusive
    hung_true = np.where(hold_i > thr, rnd > 20, ~(rnd > 10))
    hung = np.where(hung_true, 0, 1).astype(np.int64)                    # 0 -> "T", 1 -> "F"
    return {"cust": cust, "issue": issue, "tod": tod, "hold": hold_i, "hung": hung, "area": area}


def call_hangup_csv(n: int, seed: int = DEFAULT_SEED, start: int = 0):
    """CSV lines `id,customer type,area code,issue,time of day,hold time,hungup`."""
    f = call_hangup_fields(n, seed, start)
    lines = []
    for i in range(n):
        lines.append("%010d,%s,%d,%s,%s,%d,%s" % (start + i, CUST_TYPES[f["cust"][i]], f["area"][i], ISSUE_CARD[f["issue"][i]],
                                                  TOD[f["tod"][i]], f["hold"][i], "TF"[f["hung"][i]]))
    return lines


def call_hangup_encoded(n: int, seed: int = DEFAULT_SEED, start: int = 0, chunk: int = 2_000_000):
    """Bin codes exactly as the host encoder derives them from the CSV text with
    resource/call_hangup.json: 4 uint8 columns (attrs 1,3,4,5; bins 2,4,2,10) + class codes with
    class code = first-seen index over the whole file, which is fixed here as T=0, F=1 ... unless the
    first row is 'F'; callers that compare against a CSV run must use `class_names`."""
    cols = [np.empty(n, dtype=np.uint8) for _ in range(4)]
    cls = np.empty(n, dtype=np.uint8)
    for s in range(0, n, chunk):
        m = min(chunk, n - s)
        f = call_hangup_fields(m, seed, start + s)
        cols[0][s:s + m] = f["cust"]
        cols[1][s:s + m] = f["issue"]
        cols[2][s:s + m] = f["tod"]
        cols[3][s:s + m] = np.searchsorted(HOLD_POINTS, f["hold"], side="left")   # #{p_k < v}
        cls[s:s + m] = f["hung"]
    first = int(cls[0])
    class_names = ["T", "F"] if first == 0 else ["F", "T"]
    if first == 1:
        cls = (1 - cls).astype(np.uint8)
    return cols, cls, [2, 4, 2, 10], class_names


# ---- generic binned data (configs C3 / C5 and property tests) ---------------------------------------
def binned_dataset(n: int, n_bins, n_classes: int, seed: int = DEFAULT_SEED, start: int = 0, signal_attrs: int = 4,
                   noise: float = 0.15):
    """Columns of uniform bin codes; the class follows a hidden rule over the first `signal_attrs`
    attributes (parity of "upper half" indicators) with `noise` uniform label noise."""
    rows = np.arange(start, start + n, dtype=np.uint64)
    cols = []
    score = np.zeros(n, dtype=np.int64)
    for a, b in enumerate(n_bins):
        code = (uniform(seed, rows, 32 + a) * b).astype(np.int64)
        code = np.minimum(code, b - 1)
        cols.append(code.astype(np.uint16 if b > 256 else np.uint8))
        if a < signal_attrs:
            score += (code * 2 >= b).astype(np.int64) * (a + 1)
    cls = score % n_classes
    flip = uniform(seed, rows, 5) < noise
    rnd = (uniform(seed, rows, 6) * n_classes).astype(np.int64)
    cls = np.where(flip, rnd, cls).astype(np.uint8)
    return cols, cls


def binned_dataset_device(n: int, n_bins, n_classes: int, seed: int = DEFAULT_SEED, start: int = 0, signal_attrs: int = 4,
                          noise: float = 0.15, device="cuda", chunk: int = 8_000_000):
    """binned_dataset, bit for bit, generated in device memory with torch integer/fp64 tensor ops (bench.py's C3/C5
    workloads: 100 M rows x 32 attributes would take minutes on the host and 3.3 GB of upload).  Returns
    ([uint8 tensor per attribute], uint8 class tensor); attributes with more than 256 bins are not supported here."""
    import torch
    assert all(b <= 256 for b in n_bins)
    M64 = (1 << 64) - 1

    def s64(v):  # uint64 constant as the int64 with the same bits
        v &= M64
        return v - (1 << 64) if v >= (1 << 63) else v

    def lsr(z, k):  # logical shift right of int64 bit patterns
        return (z >> k) & ((1 << (64 - k)) - 1)

    def u01(rows, draw):
        x = rows * 1024 + s64(((seed << 40) & M64) + draw) + s64(0x9E3779B97F4A7C15)
        z = (x ^ lsr(x, 30)) * s64(0xBF58476D1CE4E5B9)
        z = (z ^ lsr(z, 27)) * s64(0x94D049BB133111EB)
        z = z ^ lsr(z, 31)
        return lsr(z, 11).to(torch.float64) * (1.0 / 9007199254740992.0)

    cols = [torch.empty(n, dtype=torch.uint8, device=device) for _ in n_bins]
    cls = torch.empty(n, dtype=torch.uint8, device=device)
    for s in range(0, n, chunk):
        m = min(chunk, n - s)
        rows = torch.arange(start + s, start + s + m, dtype=torch.int64, device=device)
        score = torch.zeros(m, dtype=torch.int64, device=device)
        for a, b in enumerate(n_bins):
            code = torch.clamp((u01(rows, 32 + a) * b).to(torch.int64), max=b - 1)
            cols[a][s:s + m] = code.to(torch.uint8)
            if a < signal_attrs:
                score += (code * 2 >= b).to(torch.int64) * (a + 1)
        c = score % n_classes
        flip = u01(rows, 5) < noise
        rnd = (u01(rows, 6) * n_classes).to(torch.int64)
        cls[s:s + m] = torch.where(flip, rnd, c).to(torch.uint8)
    return cols, cls
