"""Seeded synthetic inputs of the BASELINE.json configs, shared by bench.py and the tests so that the
GPU path, the oracle and the compiled reference all see the same bytes. Host-side numpy only; nothing here is timed.

Every stream is keyed by (seed, rank, column): splitmix64 over consecutive counters for the uniform columns, a Philox
generator for the Zipf ranks.
"""
import numpy as np


def splitmix64(x):
    with np.errstate(over="ignore"):
        x = x + np.uint64(0x9E3779B97F4A7C15)
        z = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def _stream(seed, rank, column, start, stop):
    base = splitmix64(np.array([seed * 4096 + rank * 8 + column], dtype=np.uint64))[0]
    with np.errstate(over="ignore"):
        return splitmix64(np.arange(start, stop, dtype=np.uint64) + base)


def uniform_doubles(seed, rank, column, rows, out=None, chunk=1 << 24):
    """doubles in [DBL_MIN, 1): never below the reference's MAX start value"""
    vals = out if out is not None else np.empty(rows, dtype=np.float64)
    tiny = np.finfo(np.float64).tiny
    for s in range(0, rows, chunk):
        e = min(rows, s + chunk)
        v = (_stream(seed, rank, column, s, e) >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)
        np.maximum(v, tiny, out=v)
        vals[s:e] = v
    return vals


def uniform_keys(seed, rank, column, rows, key_range, out=None, chunk=1 << 24, scale=1):
    keys = out if out is not None else np.empty(rows, dtype=np.int64)
    for s in range(0, rows, chunk):
        e = min(rows, s + chunk)
        keys[s:e] = (_stream(seed, rank, column, s, e) % np.uint64(key_range)).astype(np.int64) * scale
    return keys


# ---- config 3: LEFT/RIGHT/FULL SortJoin on an (int64, int32) key, Zipf(1.1)-skewed left side -------------------------
def zipf_ranks(seed, rank, rows, domain, s=1.1):
    """`rows` Zipf(s) ranks in [1, domain] (rank 1 carries ~11-12 % of the mass at s = 1.1): numpy's unbounded Zipf
    sampler with the draws beyond the domain rejected."""
    rng = np.random.Generator(np.random.Philox(key=seed * 4096 + rank))
    out = np.empty(rows, dtype=np.int64)
    have = 0
    while have < rows:
        want = int((rows - have) * 1.35) + 1024
        z = rng.zipf(s, want)
        z = z[(z >= 1) & (z <= domain)]
        take = min(len(z), rows - have)
        out[have:have + take] = z[:take]
        have += take
    return out


def sortjoin_tables(rows, world, rank, seed=3):
    """(left, right) column lists of this rank. Right: distinct tuples (k0 = global row id, k1 = k0 mod 65536). Left: k0 =
    Zipf rank - 1 over the right side's global k0 domain, k1 = k0 mod 65536, so every left row matches exactly one right
    row (possibly on another rank) and the hot tuple lands on one GPU."""
    domain = rows * world
    rk0 = np.arange(rank * rows, (rank + 1) * rows, dtype=np.int64)
    lk0 = zipf_ranks(seed, rank, rows, domain) - 1
    left = [lk0, (lk0 % 65536).astype(np.int32), uniform_doubles(seed, rank, 2, rows)]
    right = [rk0, (rk0 % 65536).astype(np.int32), uniform_doubles(seed + 1, rank, 2, rows)]
    return left, right


# ---- config 4: group-by over `groups` distinct int64 keys -------------------------------------------------------------
def groupby_table(rows, rank, groups=1_000_000, seed=4, out_keys=None, out_vals=None):
    keys = uniform_keys(seed, rank, 0, rows, groups, out=out_keys, scale=7919)  # value * 7919: not trivially dense ids
    vals = uniform_doubles(seed, rank, 1, rows, out=out_vals)
    return [keys, vals]


# ---- config 5a: multi-column tables with a string column --------------------------------------------------------------
_DICT = {}


def word_dictionary(nwords=1_000_000, seed=5):
    """(offsets int64[nwords+1], bytes uint8[...]) of `nwords` words of length U{4..24} over a-z"""
    key = (nwords, seed)
    if key not in _DICT:
        lens = (_stream(seed, 0, 6, 0, nwords) % np.uint64(21)).astype(np.int64) + 4
        offs = np.zeros(nwords + 1, dtype=np.int64)
        np.cumsum(lens, out=offs[1:])
        data = (_stream(seed, 0, 7, 0, int(offs[-1])) % np.uint64(26)).astype(np.uint8) + ord("a")
        _DICT[key] = (offs, data)
    return _DICT[key]


def setop_table(rows, rank, which, seed=5, int_range=50_000_000, nwords=1_000_000):
    """columns (int64 a, string w, double f(a)) where the string and the double are functions of the row's id, so equal
    ids give equal rows. Table `which` = 1 draws ids from [range/2, 3*range/2): half of it overlaps table 0."""
    ids = uniform_keys(seed + which, rank, 0, rows, int_range) + (int_range // 2 if which else 0)
    word = (splitmix64(ids.astype(np.uint64)) % np.uint64(nwords)).astype(np.int64)
    doffs, ddata = word_dictionary(nwords, seed)
    lens = doffs[word + 1] - doffs[word]
    offs = np.zeros(rows + 1, dtype=np.int64)
    np.cumsum(lens, out=offs[1:])
    starts = np.repeat(doffs[word] - offs[:-1], lens)
    data = ddata[np.arange(int(offs[-1]), dtype=np.int64) + starts]
    f = (ids % 1000).astype(np.float64) * 0.125 + 1.0
    return [ids, (offs.astype(np.int32), data), f]
