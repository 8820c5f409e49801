"""TEST INFRASTRUCTURE: a CPU stand-in for one shard of the sharded filter.

It implements the shard interface smc_physics_b200.sharded.ShardedFilter drives (init / extend / weight_sums /
payload_words / gather_local / pack / unpack / rejuvenate) with numpy and the oracle's integer weight arithmetic, so that the
host logic of the N > 1 path (decision, offspring ranges, chunk planning, the exchanges) runs on CPU over gloo.
The "physics" is a synthetic, deterministic log-likelihood; the payload of a particle is (identity, elasticity,
steps survived).  Never imported by the product.
"""
import numpy as np

from oracle import oracle as orc


def synthetic_loglik(slot_global, e, t):
    """Deterministic stand-in for the observation log weight of a particle at PF step t."""
    return -0.5 * ((e - 0.6) / 0.15) ** 2 * (0.6 + 0.05 * t) + 0.3 * np.sin(0.37 * slot_global + 1.3 * t)


class CpuShard:
    WORDS = 3

    def __init__(self, n, rank, world):
        self.n, self.rank, self.N = n, rank, n * world
        self.off = rank * n

    def init(self):
        g = np.arange(self.off, self.off + self.n)
        self.ident = g.astype(np.float32)                    # which prior particle this one descends from
        self.e = ((g * 0.6180339887) % 1.0).astype(np.float32)
        self.age = np.zeros(self.n, np.float32)
        self.lw = np.zeros(self.n)

    def extend(self, t, local_max):
        g = np.arange(self.off, self.off + self.n)
        self.lw = self.lw + synthetic_loglik(g, self.e.astype(np.float64), t)
        self.age += 1
        local_max[0] = float(self.lw.max())

    def weight_sums(self, global_max, sums3):
        q = orc.weight_q(self.lw, float(global_max[0]))
        self.q = q
        q19 = q >> np.uint64(20)
        vals = np.array([q.sum(dtype=np.uint64), q19.sum(dtype=np.uint64), (q19 * q19).sum(dtype=np.uint64)], np.uint64)
        sums3.copy_(__import__("torch").from_numpy(vals.view(np.int64)))

    def payload_words(self, t):
        return self.WORDS

    def _ancestors(self, global_max, W_before, W_total, u0, k_lo, k_hi):
        # like the CUDA phases (k_cdf), the CDF is rebuilt from the CURRENT log weights on every call, so a phase
        # order that resets log weights before the last pack is caught here too
        q = orc.weight_q(self.lw, float(global_max[0]))
        cdf = np.cumsum(q, dtype=np.uint64) + np.uint64(W_before)
        U0 = int(u0 * 4294967296.0)
        tau = np.array([orc.lib().oracle_sys_threshold(U0, k, W_total, self.N) for k in range(k_lo, k_hi)],
                       dtype=np.uint64)
        return np.minimum(np.searchsorted(cdf, tau, side="right"), self.n - 1)

    def _next(self):
        if getattr(self, "_nxt", None) is None:
            self._nxt = (np.empty_like(self.ident), np.empty_like(self.e), np.empty_like(self.age))
            self._filled = 0
        return self._nxt

    def gather_local(self, t, global_max, W_before, W_total, u0, k_lo, k_hi, slot0):
        """offspring [k_lo, k_hi) that stay on this shard go straight to the local slots slot0.."""
        nxt = self._next()
        cnt = k_hi - k_lo
        if cnt > 0:
            anc = self._ancestors(global_max, W_before, W_total, u0, k_lo, k_hi)
            for dst, f in zip(nxt, (self.ident, self.e, self.age)):
                dst[slot0:slot0 + cnt] = f[anc]
            self.lw[slot0:slot0 + cnt] = 0.0      # k_gather resets the (single-buffered) log weights of the slots it fills
        self._filled += cnt

    def pack(self, t, global_max, W_before, W_total, u0, k_lo, k_hi, spr, send):
        anc = self._ancestors(global_max, W_before, W_total, u0, k_lo, k_hi)
        out = send.numpy()
        fields = (self.ident, self.e, self.age)
        pos = 0
        for d in range((self.N + spr - 1) // spr):
            a, b = max(k_lo, d * spr), min(k_hi, (d + 1) * spr)
            cnt = b - a
            if cnt <= 0:
                continue
            sel = anc[a - k_lo:b - k_lo]
            for w, f in enumerate(fields):                    # chunk layout: SoA [word][count]
                out[pos + w * cnt: pos + (w + 1) * cnt] = f[sel]
            pos += cnt * self.WORDS

    def unpack(self, t, recv, counts, slot0, word_off):
        buf = recv.numpy()
        ident, e, age = self._next()
        for c, s0, wo in zip(counts, slot0, word_off):
            if c == 0:
                continue
            ident[s0:s0 + c] = buf[wo:wo + c]
            e[s0:s0 + c] = buf[wo + c:wo + 2 * c]
            age[s0:s0 + c] = buf[wo + 2 * c:wo + 3 * c]
            self._filled += c
        assert self._filled == self.n, "gather_local + unpack must fill every local slot exactly once"
        self.ident, self.e, self.age = ident, e, age
        self._nxt = None
        self.lw = np.zeros(self.n)

    def rejuvenate(self, t, R):
        pass


def serial_reference(N, T, seed, resample_offset):
    """The same synthetic filter on one shard, resampled with the oracle's systematic resampling."""
    g = np.arange(N)
    ident, e, lw = g.astype(np.float32), ((g * 0.6180339887) % 1.0).astype(np.float32), np.zeros(N)
    resampled = []
    for t in range(T):
        lw = lw + synthetic_loglik(g, e.astype(np.float64), t)
        st = orc.weight_stats(lw)
        resampled.append(st["below_half"])
        if st["below_half"]:
            idx = orc.resample_systematic(lw, resample_offset(seed, t))
            ident, e, lw = ident[idx], e[idx], np.zeros(N)
    return ident, e, lw, resampled
