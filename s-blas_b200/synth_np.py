"""numpy restatement of the device-side synthetic generators (csrc/synth.cu): same counter-based
hashes, bit-identical output.  Used by the -m gpu tests to check the CUDA generators and by
bench.py --impl reference to build inputs without touching CUDA (small sizes / row samples)."""
import numpy as np

U64 = np.uint64
MASK = (1 << 64) - 1


def splitmix64(x):
    x = (np.asarray(x, dtype=U64) + U64(0x9E3779B97F4A7C15))
    x = (x ^ (x >> U64(30))) * U64(0xBF58476D1CE4E5B9)
    x = (x ^ (x >> U64(27))) * U64(0x94D049BB133111EB)
    return x ^ (x >> U64(31))


def hash3(seed, stream, idx):
    with np.errstate(over="ignore"):
        s = U64((int(seed) ^ ((int(stream) * 0xD1342543DE82EF95) & MASK)) & MASK)
        return splitmix64(splitmix64(s) + np.asarray(idx, dtype=U64))


def u01(h, dtype):
    if np.dtype(dtype) == np.float64:
        return (h >> U64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)
    return (h >> U64(40)).astype(np.float32) * np.float32(1.0 / 16777216.0)


def fill_uniform(n, seed, stream, lo, hi, dtype):
    dt = np.dtype(dtype).type
    u = u01(hash3(seed, stream, np.arange(n, dtype=U64)), dtype)
    return (dt(lo) + (dt(hi) - dt(lo)) * u).astype(dtype)


def laplace2d(nx, ny, dtype):
    n = nx * ny
    rows, cols, vals = [], [], []
    rp = np.zeros(n + 1, np.int32)
    i = np.arange(n)
    y, x = i // nx, i % nx
    ent = [(y > 0, i - nx, -1.0), (x > 0, i - 1, -1.0), (np.ones(n, bool), i, 4.0),
           (x < nx - 1, i + 1, -1.0), (y < ny - 1, i + nx, -1.0)]
    cnt = sum(m.astype(np.int64) for m, _, _ in ent)
    rp[1:] = np.cumsum(cnt)
    ci = np.zeros(rp[-1], np.int32)
    v = np.zeros(rp[-1], dtype)
    pos = rp[:-1].astype(np.int64).copy()
    for m, c, val in ent:
        ci[pos[m]] = c[m]
        v[pos[m]] = val
        pos[m] += 1
    return rp, ci, v


def uniform(rows, cols, deg, seed, dtype):
    rp = (np.arange(rows + 1, dtype=np.int64) * deg).astype(np.int32)
    ci = np.zeros(rows * deg, np.int32)
    for r in range(rows):
        got, t = [], 0
        while len(got) < deg:
            c = int(hash3(seed, 2 * r, t)) % cols
            t += 1
            if c not in got:
                got.append(c)
        ci[r * deg:(r + 1) * deg] = sorted(got)
    v = fill_uniform(rows * deg, seed, 1, -1.0, 1.0, dtype)
    return rp, ci, v


def banded(n, bw, seed, dtype):
    rp = (np.arange(n + 1, dtype=np.int64) * bw).astype(np.int32)
    c0 = np.clip(np.arange(n) - bw // 2, 0, n - bw)
    ci = (c0[:, None] + np.arange(bw)[None, :]).astype(np.int32).reshape(-1)
    v = fill_uniform(n * bw, seed, 1, -1.0, 1.0, dtype)
    return rp, ci, v


def rmat(scale, ef, seed, dtype):
    ne = ef << scale
    e = np.arange(ne, dtype=U64)
    r = np.zeros(ne, U64)
    c = np.zeros(ne, U64)
    for l in range(scale):
        t = hash3(seed, 3 + l, e) >> U64(32)
        q = np.where(t < U64(2448131358), 0, np.where(t < U64(3264175144), 1,
                                                       np.where(t < U64(4080218931), 2, 3))).astype(U64)
        r = (r << U64(1)) | (q >> U64(1))
        c = (c << U64(1)) | (q & U64(1))
    key = (r << U64(32)) | c
    w = (hash3(seed, 2, e) >> U64(40)).astype(np.int64)
    uk, inv = np.unique(key, return_inverse=True)
    ws = np.zeros(len(uk), np.int64)
    np.add.at(ws, inv, w)
    rows = (uk >> U64(32)).astype(np.int64)
    rp = np.zeros((1 << scale) + 1, np.int32)
    rp[1:] = np.cumsum(np.bincount(rows, minlength=1 << scale))
    ci = (uk & U64(0xffffffff)).astype(np.int32)
    v = (ws.astype(np.float64) * (1.0 / 16777216.0)).astype(dtype)
    return rp, ci, v
