"""One particle filter sharded over the ranks of a torch.distributed group (one process per GPU).

The product path is INSIDE libsmcphys (include/smcphys.h, smcphys_comm_init_rank + smcphys_run_smc): NCCL all-reduce /
all-gather for the weight normalisation, the ESS decision and the offspring ranges on the device, offspring stored
into the destination rank's buffers through peer memory, no host synchronisation per step.  `ShardedFilter` below only
ships the NCCL unique id over torch.distributed and makes the same set_trial / run_smc / get_particles calls as a
single-GPU caller.  The phase-by-phase driver that follows (`PhaseShardedFilter`, `LocalShards`) is the same protocol
spelled out on the host over the smcphys_pf_* phase API; it is what the CPU tests (gloo, world_size 2 and 3) exercise
with a stand-in shard, and a second, independent implementation the device path is checked against.

Rank r holds the contiguous global particle slots [r*n, (r+1)*n).  Extend / weight / rejuvenate need no
communication.  Per PF step the ranks exchange (a) one double (MAX all-reduce of the largest log weight), (b) three
integers per rank (all-gather of the integer weight sums W, S1, S2) from which every rank takes the same ESS
decision (Gen.maybe_resample!, Inference/particle_filter.jl:55) and, when the step resamples, (c) the offspring
payloads: systematic resampling is order preserving, so the offspring of rank r's particles fill one contiguous
range of global slots and travel as at most G contiguous chunks (all_to_all_single over NCCL / NVLink).
All arithmetic that decides indices is integer (csrc/rng.cuh), so the ancestors are bit-identical to the
single-shard filter for any number of ranks.

The planning functions below are pure host code (tested on CPU with gloo, world_size 2); the per-shard compute
is the libsmcphys phase API (include/smcphys.h, smcphys_pf_*).
"""
import math

import numpy as np
import torch
import torch.distributed as dist

from . import api

_MASK64 = (1 << 64) - 1


def offspring_ranges(W_all, N, u0):
    """Global slot range [k_lo, k_hi) filled by the offspring of each rank, from the integer weight totals."""
    W_total = int(sum(W_all))
    out, before = [], 0
    for W in W_all:
        out.append(api.offspring_range(before, int(W), W_total, N, u0))
        before += int(W)
    return out, W_total


def plan_migration(ranges, spr, rank):
    """Split the exchange into chunks.  ranges[s] = offspring range of rank s; rank d owns slots [d*spr, (d+1)*spr).

    Returns (send_counts[G], recv_counts[G], recv_slot0[G]): particles this rank sends to each destination, receives
    from each source, and the local slot where each received chunk starts."""
    G = len(ranges)
    lo, hi = ranges[rank]
    send = [max(0, min(hi, (d + 1) * spr) - max(lo, d * spr)) for d in range(G)]
    my_lo, my_hi = rank * spr, (rank + 1) * spr
    recv, slot0 = [], []
    for s in range(G):
        a, b = max(ranges[s][0], my_lo), min(ranges[s][1], my_hi)
        recv.append(max(0, b - a))
        slot0.append(max(0, a - my_lo) if b > a else 0)
    return send, recv, slot0


class _CudaShard:
    """Adapter: tensors -> device pointers of the libsmcphys phase API."""

    def __init__(self, pf):
        self.pf = pf

    def init(self):
        self.pf.pf_init()

    def extend(self, t, local_max):
        self.pf.pf_extend(t, local_max.data_ptr())

    def weight_sums(self, global_max, sums3):
        self.pf.pf_weight_sums(global_max.data_ptr(), sums3.data_ptr())

    def payload_words(self, t):
        return self.pf.pf_payload_words(t)

    def pack(self, t, global_max, W_before, W_total, u0, k_lo, k_hi, spr, send):
        self.pf.pf_pack(t, global_max.data_ptr(), W_before, W_total, u0, k_lo, k_hi, spr, send.data_ptr())

    def gather_local(self, t, global_max, W_before, W_total, u0, k_lo, k_hi, slot0):
        self.pf.pf_gather_local(t, global_max.data_ptr(), W_before, W_total, u0, k_lo, k_hi, slot0)

    def unpack(self, t, recv, counts, slot0, word_off):
        self.pf.pf_unpack(t, recv.data_ptr(), counts, slot0, word_off)

    def rejuvenate(self, t, R):
        self.pf.pf_rejuvenate(t, R)


class ShardedFilter:
    """run_smc (Inference/particle_filter.jl:37-81) over particles_per_rank x world_size particles, one rank per GPU,
    all of it inside libsmcphys (NCCL + peer memory); torch.distributed only carries the 128-byte NCCL id."""

    def __init__(self, particles_per_rank, shape=api.SPHERE, substeps=5, max_timesteps=120, rejuvenation_moves=1,
                 device=0, seed=1234, stream=None, group=None, **cfg):
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.n = int(particles_per_rank)
        self.N = self.n * self.world
        if self.N > (1 << 24):
            raise ValueError("global particle count %d exceeds 2^24 (integer weight arithmetic bound)" % self.N)
        self.R = int(rejuvenation_moves)
        self.pf = api.ParticleFilter(num_particles=self.n, global_particles=self.N, particle_offset=self.rank * self.n,
                                     shape=shape, substeps=substeps, max_timesteps=max_timesteps,
                                     rejuvenation_moves=rejuvenation_moves, device=device, seed=seed, **cfg)
        if stream is not None:
            self.pf.set_stream(stream.cuda_stream)
        if self.world > 1:
            box = [api.ParticleFilter.comm_unique_id() if self.rank == 0 else None]
            dist.broadcast_object_list(box, src=0, group=group)
            self.pf.comm_init(box[0], self.world, self.rank)
        self.T = 0

    def set_trial(self, init_pos, init_vel, obs):
        self.T = np.asarray(obs).reshape(-1, 3).shape[0]
        self.pf.set_trial(init_pos, init_vel, obs)

    def run_smc(self, R=-1, want_stats=True):
        return self.pf.run_smc(self.R if R < 0 else R, want_stats=want_stats)

    def get_particles(self, **kw):
        return self.pf.get_particles(**kw)

    def get_log_ml(self):
        return self.pf.get_log_ml()

    def close(self):
        self.pf.close()


class PhaseShardedFilter:
    """The same filter driven phase by phase from the host (smcphys_pf_* + torch.distributed collectives)."""

    def __init__(self, particles_per_rank, shape=api.SPHERE, substeps=5, max_timesteps=120, rejuvenation_moves=1,
                 device=0, seed=1234, stream=None, group=None, shard=None, **cfg):
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.n = int(particles_per_rank)
        self.N = self.n * self.world
        if self.N > (1 << 24):
            raise ValueError("global particle count %d exceeds 2^24 (integer weight arithmetic bound)" % self.N)
        self.seed, self.R = int(seed), int(rejuvenation_moves)
        self.ess_num, self.ess_den = int(cfg.get("ess_num", 1)), int(cfg.get("ess_den", 2))
        if shard is None:
            self.pf = api.ParticleFilter(num_particles=self.n, global_particles=self.N,
                                         particle_offset=self.rank * self.n, shape=shape, substeps=substeps,
                                         max_timesteps=max_timesteps, rejuvenation_moves=rejuvenation_moves,
                                         device=device, seed=seed, **cfg)
            self.shard = _CudaShard(self.pf)
            self.dev = torch.device("cuda", device)
            self.stream = stream if stream is not None else torch.cuda.Stream(device=self.dev)
            self.pf.set_stream(self.stream.cuda_stream)
        else:  # CPU stand-in used by the gloo tests of the host logic
            self.pf, self.shard, self.dev, self.stream = None, shard, torch.device("cpu"), None
        self._max = torch.zeros(1, dtype=torch.float64, device=self.dev)
        self._sums = torch.zeros(3, dtype=torch.int64, device=self.dev)
        self._all = torch.zeros(3 * self.world, dtype=torch.int64, device=self.dev)
        self.T, self.log_ml, self.stats, self.last_ranges = 0, 0.0, [], None
        self.migrated_particles = 0

    # ---- trial / results
    def set_trial(self, init_pos, init_vel, obs):
        self.T = np.asarray(obs).reshape(-1, 3).shape[0]
        if self.pf is not None:
            self.pf.set_trial(init_pos, init_vel, obs)

    def get_particles(self, **kw):
        return self.pf.get_particles(**kw)

    def close(self):
        if self.pf is not None:
            self.pf.close()

    def _ctx(self):
        return torch.cuda.stream(self.stream) if self.stream is not None else _Null()

    # ---- the filter
    def step(self, t, R):
        sh, G = self.shard, self.world
        sh.extend(t, self._max)
        if G > 1:
            dist.all_reduce(self._max, op=dist.ReduceOp.MAX, group=self.group)
        sh.weight_sums(self._max, self._sums)
        if G > 1:
            dist.all_gather_into_tensor(self._all, self._sums, group=self.group)
            host = self._all.cpu().numpy()
        else:
            host = self._sums.cpu().numpy()
        sums = host.view(np.uint64).reshape(G, 3)
        W_all = [int(x) for x in sums[:, 0]]
        S1, S2 = int(sums[:, 1].sum(dtype=np.uint64)), int(sums[:, 2].sum(dtype=np.uint64))
        W_total = sum(W_all)
        gmax = float(self._max.item())
        lse = api.lib().smcphys_lse_value(gmax, W_total & _MASK64)
        resample = bool(api.lib().smcphys_ess_below(S1, S2, self.N, self.ess_num, self.ess_den))
        self.stats.append(dict(ess=api.lib().smcphys_ess_value(S1, S2), lse=lse, resampled=resample, accept_rate=None))
        if resample:
            self.log_ml += lse - math.log(self.N)
            u0 = api.lib().smcphys_resample_offset(self.seed, t)
            ranges, _ = offspring_ranges(W_all, self.N, u0)
            self.last_ranges = ranges
            send, recv, slot0 = plan_migration(ranges, self.n, self.rank)
            words = sh.payload_words(t)
            k_lo, k_hi = ranges[self.rank]
            W_before = sum(W_all[:self.rank])
            my_lo, my_hi = self.rank * self.n, (self.rank + 1) * self.n
            # only the ranges below / above the own slots travel: chunks for ranks < rank, then ranks > rank.  They are
            # packed BEFORE the local gather: k_gather resets the log weights of the slots it fills, and pack rebuilds
            # the integer CDF from the log weights.
            l_lo, l_hi = max(k_lo, my_lo), min(k_hi, my_hi)
            send[self.rank] = recv[self.rank] = 0
            below, above = (k_lo, min(k_hi, my_lo)), (max(k_lo, my_hi), k_hi)
            n_below = max(0, below[1] - below[0])
            sbuf = torch.empty(max(1, sum(send) * words), dtype=torch.float32, device=self.dev)
            rbuf = torch.empty(max(1, sum(recv) * words), dtype=torch.float32, device=self.dev)
            if n_below > 0:
                sh.pack(t, self._max, W_before, W_total, u0, below[0], below[1], self.n, sbuf)
            if above[1] > above[0]:
                sh.pack(t, self._max, W_before, W_total, u0, above[0], above[1], self.n, sbuf[n_below * words:])
            # offspring that stay on this rank (the bulk: resampling preserves order) are gathered in place
            if l_hi > l_lo:
                sh.gather_local(t, self._max, W_before, W_total, u0, l_lo, l_hi, l_lo - my_lo)
            else:
                sh.gather_local(t, self._max, W_before, W_total, u0, my_lo, my_lo, 0)
            if G > 1:
                dist.all_to_all_single(rbuf[:sum(recv) * words], sbuf[:sum(send) * words], [c * words for c in recv],
                                       [c * words for c in send], group=self.group)
            word_off = np.concatenate([[0], np.cumsum([c * words for c in recv])[:-1]]).astype(np.int64)
            sh.unpack(t, rbuf, recv, slot0, word_off)
            self.migrated_particles += sum(recv)
        if R > 0:
            sh.rejuvenate(t, R)

    def run_smc(self, R=-1, want_stats=True):
        R = self.R if R < 0 else R
        self.stats, self.log_ml, self.migrated_particles = [], 0.0, 0
        with self._ctx():
            self.shard.init()
            for t in range(self.T):
                self.step(t, R)
        return self.stats if want_stats else None


class _Null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


class LocalShards:
    """The same sharded filter with all G shards inside ONE process (shards on one device or one per device),
    exchanging through device-to-device copies instead of NCCL.  This is how a single `ccall`ing host process
    (the Julia side) can span several GPUs, and how the sharded phases are verified on a single GPU: its
    particles are bit-identical to the unsharded filter with N = G * n."""

    def __init__(self, particles_per_shard, n_shards, devices=None, shape=api.SPHERE, substeps=5, max_timesteps=120,
                 rejuvenation_moves=1, seed=1234, **cfg):
        self.n, self.G = int(particles_per_shard), int(n_shards)
        self.N, self.seed, self.R = self.n * self.G, int(seed), int(rejuvenation_moves)
        self.ess_num, self.ess_den = int(cfg.get("ess_num", 1)), int(cfg.get("ess_den", 2))
        devices = list(devices) if devices is not None else [0] * self.G
        self.devs = [torch.device("cuda", d) for d in devices]
        self.pfs = [api.ParticleFilter(num_particles=self.n, global_particles=self.N, particle_offset=r * self.n,
                                       shape=shape, substeps=substeps, max_timesteps=max_timesteps,
                                       rejuvenation_moves=rejuvenation_moves, device=devices[r], seed=seed, **cfg)
                    for r in range(self.G)]
        self.streams = [torch.cuda.Stream(device=d) for d in self.devs]
        for pf, s in zip(self.pfs, self.streams):
            pf.set_stream(s.cuda_stream)
        self._max = [torch.zeros(1, dtype=torch.float64, device=d) for d in self.devs]
        self._sums = [torch.zeros(3, dtype=torch.int64, device=d) for d in self.devs]
        self.T, self.log_ml, self.stats = 0, 0.0, []

    def set_trial(self, init_pos, init_vel, obs):
        self.T = np.asarray(obs).reshape(-1, 3).shape[0]
        for pf in self.pfs:
            pf.set_trial(init_pos, init_vel, obs)

    def close(self):
        for pf in self.pfs:
            pf.close()

    def _sync(self):
        for pf in self.pfs:
            pf.synchronize()

    def step(self, t, R):
        G, L = self.G, api.lib()
        for r in range(G):
            self.pfs[r].pf_extend(t, self._max[r].data_ptr())
        self._sync()
        gmax = max(float(m.item()) for m in self._max)
        for r in range(G):
            self._max[r].fill_(gmax)
        torch.cuda.synchronize()
        for r in range(G):
            self.pfs[r].pf_weight_sums(self._max[r].data_ptr(), self._sums[r].data_ptr())
        self._sync()
        sums = np.stack([s.cpu().numpy().view(np.uint64) for s in self._sums])
        W_all = [int(x) for x in sums[:, 0]]
        S1, S2, W_total = int(sums[:, 1].sum(dtype=np.uint64)), int(sums[:, 2].sum(dtype=np.uint64)), sum(W_all)
        lse = L.smcphys_lse_value(gmax, W_total & _MASK64)
        resample = bool(L.smcphys_ess_below(S1, S2, self.N, self.ess_num, self.ess_den))
        self.stats.append(dict(ess=L.smcphys_ess_value(S1, S2), lse=lse, resampled=resample, accept_rate=None))
        if resample:
            self.log_ml += lse - math.log(self.N)
            u0 = L.smcphys_resample_offset(self.seed, t)
            ranges, _ = offspring_ranges(W_all, self.N, u0)
            words = self.pfs[0].pf_payload_words(t)
            plans = [plan_migration(ranges, self.n, r) for r in range(G)]
            sbufs = []
            for r in range(G):
                k_lo, k_hi = ranges[r]
                sb = torch.empty(max(1, (k_hi - k_lo) * words), dtype=torch.float32, device=self.devs[r])
                self.pfs[r].pf_pack(t, self._max[r].data_ptr(), sum(W_all[:r]), W_total, u0, k_lo, k_hi, self.n,
                                    sb.data_ptr())
                sbufs.append(sb)
            self._sync()
            for d in range(G):
                send_d, recv, slot0 = plans[d]
                recv = list(recv)
                # the chunk that stays on shard d is gathered in place (as ShardedFilter does); the rest is copied
                l_lo, l_hi = max(ranges[d][0], d * self.n), min(ranges[d][1], (d + 1) * self.n)
                if l_hi <= l_lo:
                    l_lo = l_hi = d * self.n
                self.pfs[d].pf_gather_local(t, self._max[d].data_ptr(), sum(W_all[:d]), W_total, u0, l_lo, l_hi,
                                            l_lo - d * self.n)
                recv[d] = 0
                rb = torch.empty(self.n * words, dtype=torch.float32, device=self.devs[d])
                off = 0
                word_off = []
                for s in range(G):
                    word_off.append(off)
                    c = recv[s] * words
                    if c:
                        src_off = sum(plans[s][0][:d]) * words   # chunks of source s are ordered by destination
                        rb[off:off + c].copy_(sbufs[s][src_off:src_off + c])
                    off += c
                torch.cuda.synchronize()
                self.pfs[d].pf_unpack(t, rb.data_ptr(), recv, slot0, word_off)
                self.pfs[d].synchronize()
        if R > 0:
            for r in range(G):
                self.pfs[r].pf_rejuvenate(t, R)

    def run_smc(self, R=-1, want_stats=True):
        R = self.R if R < 0 else R
        self.stats, self.log_ml = [], 0.0
        for pf in self.pfs:
            pf.pf_init()
        for t in range(self.T):
            self.step(t, R)
        self._sync()
        return self.stats if want_stats else None

    def get_particles(self, **kw):
        outs = [pf.get_particles(**kw) for pf in self.pfs]
        return {k: (np.concatenate([o[k] for o in outs], axis=-1) if outs[0][k] is not None else None) for k in outs[0]}
