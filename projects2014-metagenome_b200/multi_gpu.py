"""Range-partitioned BOSS chunk build across the GPUs of one box.

One process per GPU (torch.distributed, NCCL over NVLink/NVSwitch for the plumbing). The path is
partitioned the way the reference's own chunked build partitions it -- by the leading characters
of the sort key (boss_chunk_construct.cpp:366-397, :462-525, :754-792, :917-931):

  1. every rank extracts the k-mers of its read shard and groups them by routing bucket
     (top part_bits bits of the KMerBOSS word)                         mtg_mg_extract
  2. all-to-all #1: every bucket goes to the rank that owns its range   NCCL all_to_all_single
  3. owners sort + fold their k-mers (+ counts)                          mtg_mg_collect
  3b. canonical mode: steps 1-3 handle the canonical k-mers only; the owners then emit both
      strands of their DISTINCT k-mers (mtg_mg_strands), a second (small) all-to-all routes them
      by the final ranges (planned on both strands) and the owners sort the reverse complements
      and merge (mtg_mg_merge) -- add_reverse_complements (boss_chunk_construct.cpp:179-222)
  4. target nodes x[2..K] of the owned k-mers, grouped by bucket         mtg_mg_targets
  5. all-to-all #2: targets go to the owner of that node range          NCCL all_to_all_single
  6. owners: dummy sinks = targets \\ sources, sources-1 = sources \\ targets   mtg_mg_dummies
  7. all-gather of the (few) sources-1 nodes; every rank expands them into dummy source
     k-mers and keeps the ones of its range; merge + chunk emit          mtg_mg_build
  8. the host concatenates the per-rank chunks with BOSS::Chunk::extend (boss_chunk.cpp:230-270).

The planning helpers (plan_ranges, split_sizes) and the chunk gathering are backend independent
and are covered by world_size-2 gloo tests on CPU; `VirtualBuild` runs the same phases for g
virtual ranks on ONE GPU (exchange = tensor slicing), which is how the single-GPU parity tests
check the partitioned path against the oracle.
"""
import sys

import numpy as np

from .constructor import BOSSChunkConstructor, Chunk


# ---- planning (pure host logic) ---------------------------------------------------------------

def plan_ranges(bucket_hist, world):
    """Contiguous bucket ranges [lo, hi) per rank, balancing the k-mer counts.
    bucket_hist: global (summed over ranks) k-mers per routing bucket."""
    hist = np.asarray(bucket_hist, dtype=np.float64)
    nb = len(hist)
    total = hist.sum()
    bounds = [0]
    csum = np.concatenate([[0.0], np.cumsum(hist)])
    for r in range(1, world):
        target = total * r / world
        b = int(np.searchsorted(csum, target, side="left"))
        # nearest boundary to the target, never before the previous one
        if b > 0 and abs(csum[b - 1] - target) <= abs(csum[min(b, nb)] - target):
            b -= 1
        b = max(bounds[-1], min(b, nb))
        bounds.append(b)
    bounds.append(nb)
    return [(bounds[r], bounds[r + 1]) for r in range(world)]


def split_sizes(bucket_counts, ranges):
    """Number of elements of a bucket-grouped array that go to every rank."""
    c = np.asarray(bucket_counts, dtype=np.int64)
    return [int(c[lo:hi].sum()) for lo, hi in ranges]


def route_addresses(all_counts, ranges, me, base_ptrs, key_bytes):
    """Where source rank `me` writes its k-mers (mtg_mg_route). all_counts[s][b] = k-mers of source s
    in routing bucket b; destination d's receive buffer (at base_ptrs[d]) holds one slice per
    source, in rank order. -> (bucket_owner uint32[nb], owner_dst uint64[world],
    sizes int64[source][destination])"""
    c = np.asarray(all_counts, dtype=np.int64)
    sizes = np.array([[int(c[s, lo:hi].sum()) for lo, hi in ranges] for s in range(c.shape[0])], dtype=np.int64)
    owner = np.zeros(c.shape[1], dtype=np.uint32)
    dst = np.zeros(len(ranges), dtype=np.uint64)
    for d, (lo, hi) in enumerate(ranges):
        owner[lo:hi] = d
        dst[d] = int(base_ptrs[d]) + int(sizes[:me, d].sum()) * key_bytes
    return owner, dst, sizes


def route_segments(all_counts, ranges, me, base_ptrs, key_bytes, char_shift):
    """Plan of mtg_mg_targets_route for source rank `me`. A SEGMENT is the part of a destination's
    bucket range that shares one leading character (bucket >> char_shift): a source's slice of a
    segment arrives as a sorted run. Destination d's buffer holds its segments in order, each
    segment the slices of all sources in rank order.
    -> (bucket_seg uint32[nb], seg_dst uint64[n_seg], runs[d] = run lengths at destination d in
    buffer order, sizes int64[source][destination])"""
    c = np.asarray(all_counts, dtype=np.int64)
    world, nb = c.shape[0], c.shape[1]
    bucket_seg = np.zeros(nb, dtype=np.uint32)
    seg_dst, runs, sizes = [], [], np.zeros((world, len(ranges)), dtype=np.int64)
    for d, (lo, hi) in enumerate(ranges):
        off, runs_d = 0, []
        b = lo
        while b < hi:
            e = min(hi, ((b >> char_shift) + 1) << char_shift)       # end of this character's buckets
            per_src = c[:, b:e].sum(axis=1)
            bucket_seg[b:e] = len(seg_dst)
            seg_dst.append(int(base_ptrs[d]) + (off + int(per_src[:me].sum())) * key_bytes)
            runs_d += [int(x) for x in per_src]
            off += int(per_src.sum())
            sizes[:, d] += per_src
            b = e
        runs.append(runs_d)
    if not seg_dst:
        seg_dst = [0]
    return bucket_seg, np.asarray(seg_dst, dtype=np.uint64), runs, sizes


def concat_chunks(chunks):
    """Host concatenation, exactly BOSS::Chunk::extend in rank (= key range) order."""
    full = chunks[0]
    for c in chunks[1:]:
        full.extend(c)
    return full


# ---- device buffers ---------------------------------------------------------------------------

class _CudaPtr:
    """Zero-copy torch view of library-owned device memory (__cuda_array_interface__)."""

    def __init__(self, ptr, n, words, typestr="<i8"):
        shape = (n, words) if words > 1 else (n,)
        self.__cuda_array_interface__ = {"shape": shape, "typestr": typestr, "data": (ptr, False),
                                         "version": 2, "strides": None}


def as_tensor(ptr, n, words, device):
    import torch
    if n == 0:
        return torch.empty((0, words) if words > 1 else (0,), dtype=torch.int64, device=device)
    return torch.as_tensor(_CudaPtr(ptr, n, words), device=device)


def as_counts(ptr, n, device):
    """uint32 k-mer counts of the library as an int32 tensor (bit pattern preserved); None if ptr is 0."""
    import torch
    if not ptr:
        return None
    if n == 0:
        return torch.empty((0,), dtype=torch.int32, device=device)
    return torch.as_tensor(_CudaPtr(ptr, n, 1, "<i4"), device=device)


def key_words(K, alphabet):
    return 1 if K * (2 if alphabet == "DNA" else 5) <= 64 else 2


# ---- the phases of one rank ---------------------------------------------------------------------

class RangedRank:
    """State of one rank (one GPU / one virtual rank) of the partitioned build."""

    def __init__(self, K, canonical, bpc, alphabet, device):
        import torch
        self.torch = torch
        self.device = torch.device("cuda", device)
        self.words = key_words(K, alphabet)
        # the mtg_mg_* phases consume the raw reads: no folding into runs at push time
        self.ctor = BOSSChunkConstructor.initialize(K - 1, canonical, bpc, alphabet=alphabet, device=device,
                                                    no_batching=True)
        self.nb = 1 << self.ctor.mg_part_bits()

    def extract(self):
        counts, ptr, n = self.ctor.mg_extract()
        self.n_extracted = n
        return counts, as_tensor(ptr, n, self.words, self.device)

    def count(self):
        counts, n = self.ctor.mg_count()
        self.n_extracted = n
        return counts

    def route(self, bucket_owner, owner_dst):
        self.ctor.mg_route(bucket_owner, owner_dst)

    def collect(self, keys):
        self._keep = keys   # the library copies; keep the tensor alive until the call returns
        n = keys.shape[0]
        self.n_real = self.ctor.mg_collect(keys.data_ptr() if n else 0, n)
        self._keep = None
        return self.n_real

    def accumulate(self, last):
        """super-batched build: fold the last collect() into this range's accumulated k-mers"""
        self.n_real = self.ctor.mg_accumulate(last)
        return self.n_real

    def strands(self):
        """canonical mode: (x_counts, r_counts, X, X counts, R, R counts) -- zero-copy views"""
        xc, rc, (px, pxc, nx), (pr, prc, nr) = self.ctor.mg_strands()
        return (xc, rc, as_tensor(px, nx, self.words, self.device), as_counts(pxc, nx, self.device),
                as_tensor(pr, nr, self.words, self.device), as_counts(prc, nr, self.device))

    def merge(self, X, Xc, R, Rc, r_runs=None):
        self._keep = (X, Xc, R, Rc)
        nx, nr = X.shape[0], R.shape[0]
        self.n_real = self.ctor.mg_merge(X.data_ptr() if nx else 0, Xc.data_ptr() if (Xc is not None and nx) else 0, nx,
                                         R.data_ptr() if nr else 0, Rc.data_ptr() if (Rc is not None and nr) else 0, nr,
                                         r_runs)
        self._keep = None
        return self.n_real

    def targets_count(self):
        counts, n = self.ctor.mg_targets_count()
        return counts

    def targets_route(self, bucket_seg, seg_dst):
        self.ctor.mg_targets_route(bucket_seg, seg_dst)

    def char_shift(self):
        """routing-bucket bits below the leading character of a node word, or None if a bucket does
        not resolve the character (then the targets are exchanged as grouped, folded slices)"""
        pb, L = self.ctor.mg_part_bits(), self.ctor.tight_bits()
        return pb - L if pb >= L else None

    def targets(self):
        counts, ptr, n = self.ctor.mg_targets()
        return counts, as_tensor(ptr, n, self.words, self.device)

    def dummies(self, targets_in, runs=None, runs_per_set=0):
        n = targets_in.shape[0]
        ptr, n_src, n_sink = self.ctor.mg_dummies(targets_in.data_ptr() if n else 0, n, runs, runs_per_set)
        self.n_sink = n_sink
        return as_tensor(ptr, n_src, self.words, self.device)

    def build(self, all_src, lo, hi, first, device_output=False):
        n = all_src.shape[0]
        return self.ctor.mg_build(all_src.data_ptr() if n else 0, n, lo, hi, first, device_output)

    def close(self):
        self.ctor.close()


def _bucket_offsets(counts, ranges):
    c = np.concatenate([[0], np.cumsum(np.asarray(counts, dtype=np.int64))])
    return [(int(c[lo]), int(c[hi])) for lo, hi in ranges]


class VirtualBuild:
    """g virtual ranks on one GPU: the same phases, the exchanges done by slicing."""

    def __init__(self, world, K, canonical=False, bits_per_count=0, alphabet="DNA", device=0, fused_route=True):
        self.world = world
        self.canonical = canonical
        self.fused_route = fused_route
        self.ranks = [RangedRank(K, canonical, bits_per_count, alphabet, device) for _ in range(world)]

    def add_sequences(self, rank, sequences):
        self.ranks[rank].ctor.add_sequences(sequences)

    def build_chunks(self, more_rounds=None):
        """more_rounds: [{rank: reads}, ...] -- further slices of the input, each pushed, routed with
        the ranges of the first round, collected and folded into the ranks' accumulated k-mers
        (super-batched build); None: everything was pushed before the call."""
        rounds = [None] + list(more_rounds or [])
        for i, extra in enumerate(rounds):
            if extra is not None:
                for r_, seqs in extra.items():
                    self.ranks[r_].ctor.add_sequences(seqs)
            self._collect_round(plan=(i == 0))
            if len(rounds) > 1:
                for r in self.ranks:
                    r.accumulate(i == len(rounds) - 1)
        return self._finish()

    def _collect_round(self, plan):
        torch = self.ranks[0].torch
        g = self.world
        if self.fused_route:
            # grouping + exchange in one kernel per rank: buckets are scattered straight into the
            # owners' receive buffers (here: plain device tensors, over NVLink: peer-mapped memory)
            counts = np.stack([r.count() for r in self.ranks])
            if plan:
                self.ranges = plan_ranges(counts.sum(axis=0), g)
            words = self.ranks[0].words
            _, _, sizes = route_addresses(counts, self.ranges, 0, [0] * g, 8 * words)
            routed = [torch.empty((int(sizes[:, d].sum()), words) if words > 1 else (int(sizes[:, d].sum()),),
                                  dtype=torch.int64, device=self.ranks[0].device) for d in range(g)]
            torch.cuda.synchronize()
            for s_, r in enumerate(self.ranks):
                owner, dst, _ = route_addresses(counts, self.ranges, s_, [t.data_ptr() for t in routed], 8 * words)
                r.route(owner, dst)
        else:
            ex = [r.extract() for r in self.ranks]
            hist = np.sum([c for c, _ in ex], axis=0)
            if plan:
                self.ranges = plan_ranges(hist, g)
            offs = [_bucket_offsets(c, self.ranges) for c, _ in ex]
            routed = [torch.cat([ex[s][1][offs[s][d][0]:offs[s][d][1]] for s in range(g)]) for d in range(g)]
        torch.cuda.synchronize()   # the library works on its own streams
        for d, r in enumerate(self.ranks):
            r.collect(routed[d])
        del routed

    def _finish(self):
        torch = self.ranks[0].torch
        g = self.world
        if self.canonical:
            # second strand from the distinct canonical k-mers, routed by the FINAL ranges
            st = [r.strands() for r in self.ranks]
            hist = np.sum([x + rr for x, rr, *_ in st], axis=0)
            self.ranges = plan_ranges(hist, g)
            xo = [_bucket_offsets(x, self.ranges) for x, *_ in st]
            ro = [_bucket_offsets(rr, self.ranges) for _, rr, *_ in st]
            counting = st[0][3] is not None

            def route(idx, offs, d):
                return torch.cat([st[s][idx][offs[s][d][0]:offs[s][d][1]] for s in range(g)])

            recv = [(route(2, xo, d), route(3, xo, d) if counting else None,
                     route(4, ro, d), route(5, ro, d) if counting else None) for d in range(g)]
            torch.cuda.synchronize()
            for d, r in enumerate(self.ranks):
                r.merge(*recv[d], r_runs=[ro[s][d][1] - ro[s][d][0] for s in range(g)])
            del recv, st
        csh = self.ranks[0].char_shift()
        if self.fused_route and csh is not None:
            tcounts = np.stack([r.targets_count() for r in self.ranks])
            words = self.ranks[0].words
            _, _, runs, sizes = route_segments(tcounts, self.ranges, 0, [0] * g, 8 * words, csh)
            recv = [torch.empty((int(sizes[:, d].sum()), words) if words > 1 else (int(sizes[:, d].sum()),),
                                dtype=torch.int64, device=self.ranks[0].device) for d in range(g)]
            torch.cuda.synchronize()
            for s_, r in enumerate(self.ranks):
                seg, dst, _, _ = route_segments(tcounts, self.ranges, s_, [t.data_ptr() for t in recv], 8 * words, csh)
                r.targets_route(seg, dst)
            per_set = g
        else:
            per_set = 0
            tg = [r.targets() for r in self.ranks]
            toffs = [_bucket_offsets(c, self.ranges) for c, _ in tg]
            recv = [torch.cat([tg[s][1][toffs[s][d][0]:toffs[s][d][1]] for s in range(g)]) for d in range(g)]
            runs = [[toffs[s][d][1] - toffs[s][d][0] for s in range(g)] for d in range(g)]
        torch.cuda.synchronize()
        src = [r.dummies(recv[d], runs[d], per_set).clone() for d, r in enumerate(self.ranks)]
        all_src = torch.cat(src)
        torch.cuda.synchronize()
        chunks = [r.build(all_src, self.ranges[d][0], self.ranges[d][1], d == 0)
                  for d, r in enumerate(self.ranks)]
        return chunks

    def build_chunk(self, more_rounds=None):
        return concat_chunks(self.build_chunks(more_rounds))

    def close(self):
        for r in self.ranks:
            r.close()


# ---- one process per GPU ------------------------------------------------------------------------

# Symmetric (peer-mapped) receive buffers live as long as the process: keyed by (group, device,
# slot), shared by every DistBuild of the process and only ever replaced by a larger one.
_SYMM_SLOTS = {}
_SYMM_RETIRED = []


def release_symmetric_buffers():
    """Drop the process-wide symmetric buffers (call on every rank before destroy_process_group)."""
    _SYMM_SLOTS.clear()
    _SYMM_RETIRED.clear()


class ExchangeMismatch(RuntimeError):
    """An exchange delivered other rows than the ones that were sent (MTG_MG_VERIFY)."""


def row_sums(t):
    """int64 wrap-around checksum of every row of a 1-D / 2-D integer tensor"""
    import torch
    if t.dim() == 1:
        return t.to(torch.int64)
    return t.to(torch.int64).sum(dim=1)


def segment_sums(t, sizes):
    """checksums of consecutive row segments of t (lengths `sizes`) -> int64[len(sizes)] on t's device"""
    import torch
    cs = torch.zeros(t.shape[0] + 1, dtype=torch.int64, device=t.device)
    if t.shape[0]:
        cs[1:] = torch.cumsum(row_sums(t), 0)
    ends = torch.cumsum(torch.as_tensor([int(x) for x in sizes], dtype=torch.int64), 0).to(t.device)
    starts = ends - torch.as_tensor([int(x) for x in sizes], dtype=torch.int64).to(t.device)
    return cs[ends] - cs[starts]


def routing_buckets(keys, words, shift, part_bits):
    """routing bucket (bits [shift, shift + part_bits) of the k-mer word) of every row of a device array
    of 64-bit (words = 1) or 128-bit (words = 2: low word first) keys -- the digit of the ROUTE scatter"""
    import torch
    if part_bits == 0:
        return torch.zeros(keys.shape[0], dtype=torch.int64, device=keys.device)
    mask = (1 << part_bits) - 1
    if words == 1:
        return (keys >> shift) & mask
    lo, hi = keys[:, 0], keys[:, 1]
    if shift >= 64:
        return (hi >> (shift - 64)) & mask
    return (((lo >> shift) & ((1 << (64 - shift)) - 1)) | (hi << (64 - shift))) & mask


class DistBuild:
    """The partitioned build of THIS rank inside a torch.distributed (NCCL) group.

    All library work of the rank runs on the torch stream that is current when the object is
    created (mtg_set_stream): the NCCL collectives, the symmetric-memory barriers and the peer
    copies issued here are ordered against the kernels of the mtg_mg_* phases through that stream.
    verify=True (or MTG_MG_VERIFY=1): after every exchange the receiver checks, per source rank,
    the number of rows and a 64-bit checksum against what the sender addressed to it, and all ranks
    raise ExchangeMismatch together if any slice differs."""

    def __init__(self, K, canonical=False, bits_per_count=0, alphabet="DNA", device=0, group=None, verify=None):
        import torch
        import torch.distributed as dist
        self.torch, self.dist, self.group = torch, dist, group
        self.world = dist.get_world_size(group)
        self.rank_id = dist.get_rank(group)
        self.rank = RangedRank(K, canonical, bits_per_count, alphabet, device)
        self.canonical = canonical
        self.K = K
        self.key_bits = K * (2 if alphabet == "DNA" else 5)
        if self.rank.device.type == "cuda":
            with torch.cuda.device(self.rank.device):
                self.stream = torch.cuda.current_stream()
            self.rank.ctor.set_stream(self.stream.cuda_stream)
        # MTG_A2A=nccl forces the NCCL all-to-all; default: peer stores over NVLink (symmetric memory)
        import os
        self.use_symm = (os.environ.get("MTG_A2A", "peer") != "nccl" and dist.get_backend(group) == "nccl"
                         and self.world > 1)
        # MTG_A2A=kernel: the pushes as ONE kernel of peer stores (csrc/mtg_peer.cu) instead of one
        # copy-engine peer copy per destination (measured at 2 GPUs: 470 vs 661 GB/s, so not the default)
        self.push_kernel = os.environ.get("MTG_A2A", "peer") == "kernel"
        # MTG_ROUTE=0: grouping pass, then a separate exchange (the first design)
        self.fused_route = os.environ.get("MTG_ROUTE", "1") != "0"
        self._slots = {}
        self._legacy_slots = os.environ.get("MTG_SYMM_LEGACY", "0") == "1"
        self._streams = [torch.cuda.Stream(device=self.rank.device) for _ in range(self.world)] if self.use_symm else []
        if self.use_symm:
            try:
                self._symm_slot(0, 1 << 20)
            except Exception as e:   # no peer mapping on this system: NCCL
                print("mtg_b200: symmetric memory unavailable (%s), using NCCL all_to_all" % e, file=sys.stderr)
                self.use_symm = False
        self.a2a_bytes = 0
        self.a2a_ms = 0.0
        self.phase_events = None
        self.verify = (os.environ.get("MTG_MG_VERIFY", "0") != "0") if verify is None else bool(verify)
        self.verified = []      # names of the exchanges checked by the last build

    # ---- verification of an exchange (debug / parity runs) --------------------------------------
    def _verify(self, what, sent_sums, sent_counts, got_sums, got_counts, column=None):
        """sent_*: int64[n] device tensors, one entry per destination slot this rank addressed;
        got_*: int64[world] (or [n_mine, world]) of what arrived here per source. `column` = the
        indices of MY slots in the senders' vectors (default: my rank). All ranks raise together."""
        torch, dist = self.torch, self.dist
        dev = self.rank.device
        mine = torch.stack([sent_sums.to(torch.int64), sent_counts.to(torch.int64)]).contiguous()   # [2, n]
        allv = torch.empty((self.world,) + tuple(mine.shape), dtype=torch.int64, device=dev)
        dist.all_gather_into_tensor(allv.view(self.world * 2, -1), mine, group=self.group)
        cols = [self.rank_id] if column is None else list(column)
        idx = torch.as_tensor(cols, dtype=torch.int64, device=dev)
        exp_sums = allv[:, 0, :].index_select(1, idx).t()      # [my slots, source]
        exp_counts = allv[:, 1, :].index_select(1, idx).t()
        got_sums = got_sums.to(dev).view(len(cols), self.world)
        got_counts = torch.as_tensor(got_counts, dtype=torch.int64).to(dev).view(len(cols), self.world)
        bad = (exp_sums != got_sums) | (exp_counts != got_counts)
        n_bad = bad.sum().view(1)
        dist.all_reduce(n_bad, group=self.group)
        if int(n_bad):
            for j, s_ in bad.nonzero().tolist():
                print("mtg_b200 VERIFY %s: rank %d slot %d: rows of source rank %d differ: expected n=%d sum=%d, "
                      "received n=%d sum=%d" % (what, self.rank_id, cols[j], s_, int(exp_counts[j, s_]),
                                                int(exp_sums[j, s_]), int(got_counts[j, s_]), int(got_sums[j, s_])),
                      file=sys.stderr, flush=True)
            raise ExchangeMismatch("exchange '%s': %d (source, destination) slices differ from what was sent"
                                   % (what, int(n_bad)))
        self.verified.append(what)

    def _pending_sums(self, which, bucket_slot, n_slots, shift):
        """checksum / number of the rows the next route call will store into every destination slot"""
        torch = self.torch
        ptr, n = self.rank.ctor.mg_debug_pending(which)
        keys = as_tensor(ptr, n, self.rank.words, self.rank.device)
        pb = self.rank.ctor.mg_part_bits()
        slot = torch.as_tensor(np.asarray(bucket_slot, dtype=np.int64), device=self.rank.device)[
            routing_buckets(keys, self.rank.words, shift, pb)]
        sums = torch.zeros(n_slots, dtype=torch.int64, device=self.rank.device).scatter_add_(0, slot, row_sums(keys))
        counts = torch.bincount(slot, minlength=n_slots)
        return sums, counts

    # ---- exchange: NVLink peer stores into symmetric memory, NCCL all_to_all as the fallback ----
    def _symm_slot(self, slot, need_bytes):
        """Symmetric (peer-mapped) receive buffer `slot` of at least need_bytes on EVERY rank; the
        decision to (re)allocate is taken from numbers all ranks share, so the rendezvous is collective."""
        import torch.distributed._symmetric_memory as symm
        name = self.group.group_name if self.group is not None else self.dist.group.WORLD.group_name
        if self._legacy_slots:
            # round-1 behaviour, kept ONLY to reproduce its 8-GPU parity miss under the exchange verifier
            # (MTG_SYMM_LEGACY=1): every DistBuild allocates, rendezvouses and frees its own buffers
            cur = self._slots.get(slot)
            if cur is None or cur[0].numel() < need_bytes:
                cap = (int(need_bytes * 1.25) + (1 << 20)) & ~0xFFF
                self._slots.pop(slot, None)
                buf = symm.empty(cap, dtype=self.torch.uint8, device=self.rank.device)
                cur = (buf, symm.rendezvous(buf, name))
                self._slots[slot] = cur
            return cur
        key = (name, str(self.rank.device), slot)
        cur = _SYMM_SLOTS.get(key)
        if cur is None or cur[0].numel() < need_bytes:
            cap = (int(need_bytes * 1.25) + (1 << 20)) & ~0xFFF
            if cur is not None:
                # An outgrown block is parked, never freed: a block that went through a rendezvous must
                # not come back from the allocator for ANOTHER slot while peers still map it as the old one
                _SYMM_RETIRED.append(cur)
            buf = symm.empty(cap, dtype=self.torch.uint8, device=self.rank.device)
            hdl = symm.rendezvous(buf, name)
            # A fresh block (and its signal pad) is initialised asynchronously by its owner: no peer may
            # signal or store into it before every rank has drained that work. Round 1 re-allocated the
            # buffers per build and went straight on to the barrier kernel: exchanges lost rows and the
            # barrier eventually hung (profiles/r02_dist_check_n8_plain_legacy_FAILS.log). Blocks are now
            # allocated once per process, and this (rare) path is fenced on the host.
            self.torch.cuda.synchronize(self.rank.device)
            self.dist.barrier(group=self.group)
            cur = (buf, hdl)
            _SYMM_SLOTS[key] = cur
        self._slots[slot] = cur
        return cur

    def _exchange(self, send, send_sizes, slot=0, what="exchange"):
        """all-to-all of a bucket-grouped tensor; send_sizes[d] rows go to rank d. Returns the rows
        received, ordered by source rank."""
        recv, sizes = self._exchange_many([(send, send_sizes, slot)], what)[0]
        self.last_recv_sizes = sizes
        return recv

    def _exchange_many(self, items, what="exchange"):
        """Several all-to-alls behind ONE size all-gather and ONE barrier: items = [(send tensor,
        send_sizes per destination, symmetric slot)]. Returns [(rows received in source-rank order,
        rows per source)]."""
        torch, dist = self.torch, self.dist
        dev = self.rank.device
        ni = len(items)
        sizes = torch.tensor([list(ss) for _, ss, _ in items], dtype=torch.int64, device=dev)   # [item][dest]
        all_sizes = torch.empty((self.world * ni, self.world), dtype=torch.int64, device=dev)   # concatenated form
        dist.all_gather_into_tensor(all_sizes, sizes, group=self.group)   # also orders the ranks: every
        all_sizes = all_sizes.cpu().numpy().reshape(self.world, ni, self.world)   # peer has consumed its slots
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        out = []
        if not self.use_symm:
            ev0.record()
            for i, (send, send_sizes, _) in enumerate(items):
                recv_sizes = [int(x) for x in all_sizes[:, i, self.rank_id]]
                recv = torch.empty((int(sum(recv_sizes)),) + tuple(send.shape[1:]), dtype=send.dtype, device=dev)
                dist.all_to_all_single(recv, send.contiguous(), recv_sizes, [int(x) for x in send_sizes],
                                       group=self.group)
                out.append((recv, recv_sizes))
                self.a2a_bytes += (int(sum(send_sizes)) - int(send_sizes[self.rank_id])) * send[0:1].numel() \
                    * send.element_size() if send.shape[0] else 0
            ev1.record()
            ev1.synchronize()
            self.a2a_ms += ev0.elapsed_time(ev1)
            self._verify_items(items, out, what)
            return out
        cur = torch.cuda.current_stream()
        plans = []
        for i, (send, send_sizes, slot) in enumerate(items):
            tail = tuple(send.shape[1:])
            row = 1
            for t in tail:
                row *= int(t)
            esz = send.element_size() * row
            need = int(all_sizes[:, i, :].sum(axis=0).max()) * esz
            buf, hdl = self._symm_slot(slot, max(need, 16))
            plans.append((send.contiguous(), [int(x) for x in send_sizes], tail, row, esz, hdl))
            self.a2a_bytes += (int(sum(send_sizes)) - int(send_sizes[self.rank_id])) * esz
        ev0.record()
        if self.push_kernel:
            # one kernel of peer stores (csrc/mtg_peer.cu), every destination served concurrently
            import ctypes
            dsts, srcs, nby = [], [], []
            for i, (send, send_sizes, tail, row, esz, hdl) in enumerate(plans):
                so = np.concatenate([[0], np.cumsum(send_sizes)])
                for k in range(self.world):
                    d = (self.rank_id + k) % self.world
                    if send_sizes[d] == 0:
                        continue
                    off = int(all_sizes[:self.rank_id, i, d].sum()) * esz     # my rows' place in d's buffer
                    dsts.append(int(hdl.buffer_ptrs[d]) + off)
                    srcs.append(send.data_ptr() + int(so[d]) * esz)
                    nby.append(send_sizes[d] * esz)
            if dsts:
                n = len(dsts)
                rc = self.rank.ctor._lib.mtg_peer_push(
                    ctypes.c_void_p(cur.cuda_stream), n, (ctypes.c_void_p * n)(*dsts),
                    (ctypes.c_void_p * n)(*srcs), (ctypes.c_uint64 * n)(*nby))
                if rc != 0:
                    raise RuntimeError("mtg_peer_push failed (%d)" % rc)
        else:
            for st in self._streams:
                st.wait_stream(cur)
            for i, (send, send_sizes, tail, row, esz, hdl) in enumerate(plans):
                so = np.concatenate([[0], np.cumsum(send_sizes)])
                for k in range(self.world):
                    d = (self.rank_id + k) % self.world          # staggered: no peer is hit by all at once
                    n_d = send_sizes[d]
                    if n_d == 0:
                        continue
                    off = int(all_sizes[:self.rank_id, i, d].sum()) * row     # my rows' place in d's buffer
                    dst = hdl.get_buffer(d, (n_d,) + tail, send.dtype, off)
                    with torch.cuda.stream(self._streams[k]):
                        dst.copy_(send[int(so[d]):int(so[d + 1])], non_blocking=True)
            for st in self._streams:
                cur.wait_stream(st)
        plans[0][5].barrier()
        ev1.record()
        ev1.synchronize()
        self.a2a_ms += ev0.elapsed_time(ev1)
        for i, (send, send_sizes, tail, row, esz, hdl) in enumerate(plans):
            recv_sizes = [int(x) for x in all_sizes[:, i, self.rank_id]]
            out.append((hdl.get_buffer(self.rank_id, (int(sum(recv_sizes)),) + tail, send.dtype, 0), recv_sizes))
        self._verify_items(items, out, what)
        return out

    def _verify_items(self, items, out, what):
        if not self.verify:
            return
        torch = self.torch
        for i, ((send, send_sizes, _), (recv, recv_sizes)) in enumerate(zip(items, out)):
            self._verify("%s item %d" % (what, i), segment_sums(send, send_sizes),
                         torch.as_tensor([int(x) for x in send_sizes], dtype=torch.int64, device=self.rank.device),
                         segment_sums(recv, recv_sizes), recv_sizes)

    def _route_kmers(self, plan=True):
        """Exchange #1 fused into the grouping pass: after the per-bucket counts of all ranks are
        known, ONE scatter kernel per rank writes every bucket straight into its owner's symmetric
        receive buffer over NVLink (mtg_mg_route). Returns the k-mers received, sources in rank order."""
        torch, dist = self.torch, self.dist
        dev = self.rank.device
        counts = self.rank.count()
        self._mark("extract+count")
        mine = torch.from_numpy(counts.astype(np.int64)).to(dev)
        allc = torch.empty((self.world, mine.numel()), dtype=torch.int64, device=dev)
        dist.all_gather_into_tensor(allc, mine, group=self.group)   # also: every peer is done with slot 0
        allc = allc.cpu().numpy()
        if plan:       # later rounds of a super-batched build keep the ranges of the first
            self.ranges = plan_ranges(allc.sum(axis=0), self.world)
        kb = 8 * self.rank.words
        _, _, sizes = route_addresses(allc, self.ranges, self.rank_id, [0] * self.world, kb)
        need = int(sizes.sum(axis=0).max()) * kb
        buf, hdl = self._symm_slot(0, max(need, 16))
        owner, dst, _ = route_addresses(allc, self.ranges, self.rank_id, [int(p) for p in hdl.buffer_ptrs], kb)
        if self.verify:
            sent = self._pending_sums(0, owner, self.world, self.key_bits - self.rank.ctor.mg_part_bits())
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        self.rank.route(owner, dst)
        hdl.barrier()
        ev1.record()
        ev1.synchronize()
        self.a2a_ms += ev0.elapsed_time(ev1)
        self.a2a_bytes += int(sizes[self.rank_id].sum() - sizes[self.rank_id, self.rank_id]) * kb
        self.last_recv_sizes = [int(x) for x in sizes[:, self.rank_id]]
        total = int(sizes[:, self.rank_id].sum())
        shape = (total, self.rank.words) if self.rank.words > 1 else (total,)
        recv = hdl.get_buffer(self.rank_id, shape, torch.int64, 0)
        if self.verify:
            self._verify("k-mers (fused route)", sent[0], sent[1], segment_sums(recv, self.last_recv_sizes),
                         self.last_recv_sizes)
        self._mark("route (group + exchange)")
        return recv

    def _route_targets(self, char_shift):
        """Exchange #3 fused into the scatter pass of the targets (mtg_mg_targets_route): every
        (source, owner, leading character) slice lands as a sorted run in the owner's buffer."""
        torch, dist = self.torch, self.dist
        dev = self.rank.device
        counts = self.rank.targets_count()
        self._mark("targets (count)")
        mine = torch.from_numpy(counts.astype(np.int64)).to(dev)
        allc = torch.empty((self.world, mine.numel()), dtype=torch.int64, device=dev)
        dist.all_gather_into_tensor(allc, mine, group=self.group)   # also: every peer is done with slot 5
        allc = allc.cpu().numpy()
        kb = 8 * self.rank.words
        _, _, _, sizes = route_segments(allc, self.ranges, self.rank_id, [0] * self.world, kb, char_shift)
        need = int(sizes.sum(axis=0).max()) * kb
        buf, hdl = self._symm_slot(5, max(need, 16))
        seg, dst, runs, _ = route_segments(allc, self.ranges, self.rank_id, [int(p) for p in hdl.buffer_ptrs], kb,
                                           char_shift)
        if self.verify:
            L = self.rank.ctor.tight_bits()
            sent = self._pending_sums(1, seg, len(dst), self.key_bits - L - self.rank.ctor.mg_part_bits())
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        self.rank.targets_route(seg, dst)
        hdl.barrier()
        ev1.record()
        ev1.synchronize()
        self.a2a_ms += ev0.elapsed_time(ev1)
        self.a2a_bytes += int(sizes[self.rank_id].sum() - sizes[self.rank_id, self.rank_id]) * kb
        total = int(sizes[:, self.rank_id].sum())
        shape = (total, self.rank.words) if self.rank.words > 1 else (total,)
        tin = hdl.get_buffer(self.rank_id, shape, torch.int64, 0)
        if self.verify:
            seg0 = sum(len(r) for r in runs[:self.rank_id]) // self.world      # my first segment (global index)
            mine = list(range(seg0, seg0 + len(runs[self.rank_id]) // self.world))
            if mine:
                self._verify("targets (fused route)", sent[0], sent[1], segment_sums(tin, runs[self.rank_id]),
                             runs[self.rank_id], column=mine)
            else:   # a rank without a segment still takes part in the collectives
                self._verify("targets (fused route)", sent[0], sent[1],
                             torch.zeros(0, dtype=torch.int64, device=dev), [], column=[])
        self._mark("route targets (scatter + exchange)")
        return tin, runs[self.rank_id]

    def _drain(self):
        """Host-synchronise the torch stream before a library phase that READS tensors torch produced
        (torch.cat, NCCL results, peer copies). The library's phases end with a stream synchronise, so
        library -> torch needs nothing; torch -> library does whenever the two run on different streams,
        and they do by default: torch's default stream has the handle 0, for which mtg_set_stream keeps
        the library's own (non-blocking) stream. The missing drain before mtg_mg_build was round 1's
        8-GPU parity miss: the expansion read the gathered source nodes before torch.cat had written them
        (DESIGN.md 5)."""
        self.torch.cuda.current_stream(self.rank.device).synchronize()

    def _mark(self, name):
        """phase timing (CUDA events on the current stream, which the library shares in bench.py)"""
        if getattr(self, "phase_events", None) is not None:
            ev = self.torch.cuda.Event(enable_timing=True)
            ev.record()
            self.phase_events.append((name, ev))

    def phase_ms(self):
        evs = self.phase_events or []
        out = {}
        for (_, a), (name, b) in zip(evs, evs[1:]):
            b.synchronize()
            out[name] = out.get(name, 0.0) + a.elapsed_time(b)
        return out

    def build(self, device_output=False, batches=None):
        """batches: optional list of callables, one per slice of this rank's reads; each is called with
        the rank's constructor and pushes its slice (super-batched build: every slice is routed with the
        ranges of the first round, collected and folded into the accumulated k-mers of the range --
        the same number of slices on every rank). None: the reads were pushed before the call."""
        torch, dist = self.torch, self.dist
        dev = self.rank.device
        self.verified = []
        self._mark("start")
        rounds = list(batches) if batches else [None]
        for i, push in enumerate(rounds):
            if push is not None:
                push(self.rank.ctor)
            if self.use_symm and self.fused_route:
                recv = self._route_kmers(plan=(i == 0))
            else:
                counts, keys = self.rank.extract()
                self._mark("extract+group")
                if i == 0:
                    hist = torch.from_numpy(counts.astype(np.int64)).to(dev)
                    dist.all_reduce(hist, group=self.group)
                    self.ranges = plan_ranges(hist.cpu().numpy(), self.world)
                recv = self._exchange(keys, split_sizes(counts, self.ranges), what="k-mers")
                self._mark("exchange kmers")
            self._drain()
            self.rank.collect(recv)
            del recv
            self._mark("collect (sort+fold)")
            if len(rounds) > 1:
                self.rank.accumulate(i == len(rounds) - 1)
                self._mark("accumulate (merge+fold)")
        if self.canonical:
            # second strand from the distinct canonical k-mers, routed by the FINAL ranges
            xc, rc, X, Xc, R, Rc = self.rank.strands()
            self._mark("strands")
            hist = torch.from_numpy((xc + rc).astype(np.int64)).to(dev)
            dist.all_reduce(hist, group=self.group)
            self.ranges = plan_ranges(hist.cpu().numpy(), self.world)
            xs, rs = split_sizes(xc, self.ranges), split_sizes(rc, self.ranges)
            # X is received in rank order = increasing stage-1 key ranges; R as one sorted run per source
            items = [(X, xs, 1), (R, rs, 2)]
            if Xc is not None:
                items += [(Xc, xs, 3), (Rc, rs, 4)]
            got = self._exchange_many(items, "strands (X, R, X counts, R counts)")
            rX, rR, r_runs = got[0][0], got[1][0], got[1][1]
            rXc, rRc = (got[2][0], got[3][0]) if Xc is not None else (None, None)
            self._mark("exchange strands")
            self._drain()
            self.rank.merge(rX, rXc, rR, rRc, r_runs)
            del rX, rR, rXc, rRc
            self._mark("merge strands")
        csh = self.rank.char_shift()
        if self.use_symm and self.fused_route and csh is not None:
            tin, runs = self._route_targets(csh)
            per_set = self.world                             # one set of `world` runs per leading character
        else:
            per_set = 0
            tcounts, tg = self.rank.targets()
            self._mark("targets")
            tin = self._exchange(tg, split_sizes(tcounts, self.ranges), what="targets")
            runs = self.last_recv_sizes                      # one sorted run per source rank
            self._mark("exchange targets")
        self._drain()
        src = self.rank.dummies(tin, runs, per_set)
        self._mark("dummies")
        n_src = torch.tensor([src.shape[0]], dtype=torch.int64, device=dev)
        all_n = torch.empty(self.world, dtype=torch.int64, device=dev)
        dist.all_gather_into_tensor(all_n, n_src, group=self.group)
        all_n = all_n.tolist()
        parts = [torch.empty((n,) + tuple(src.shape[1:]), dtype=torch.int64, device=dev) for n in all_n]
        self._all_gather_ragged(parts, src, all_n)
        all_src = torch.cat(parts)
        if self.verify:
            # every rank must hold the same list: compare the per-source checksums with the owners' own
            mine = torch.zeros(self.world, dtype=torch.int64, device=dev)
            mine[self.rank_id] = row_sums(src).sum()
            cnt = torch.zeros(self.world, dtype=torch.int64, device=dev)
            cnt[self.rank_id] = src.shape[0]
            dist.all_reduce(mine, group=self.group)
            dist.all_reduce(cnt, group=self.group)
            got = segment_sums(all_src, all_n)
            bad = ((got != mine) | (cnt != torch.as_tensor(all_n, dtype=torch.int64, device=dev))).sum().view(1)
            dist.all_reduce(bad, group=self.group)
            if int(bad):
                raise ExchangeMismatch("gathered source nodes differ from the owners' lists")
            self.verified.append("source nodes (all-gather)")
        lo, hi = self.ranges[self.rank_id]
        self._mark("gather sources")
        self._drain()      # all_src = torch.cat(NCCL broadcasts) must be written before the library copies it
        out = self.rank.build(all_src, lo, hi, self.rank_id == 0, device_output)
        self._mark("build (expand+merge+emit)")
        return out

    def _all_gather_ragged(self, parts, src, all_n):
        dist = self.dist
        for r in range(self.world):   # broadcast per rank: the lists are tiny
            buf = src.contiguous() if r == self.rank_id else parts[r]
            if all_n[r]:
                dist.broadcast(buf, src=dist.get_global_rank(self.group, r) if self.group else r,
                               group=self.group)
            parts[r] = buf

    def close(self):
        self._slots.clear()
        self.rank.close()


def gather_chunks(chunk, dist, group=None, dst=0):
    """Per-rank chunks -> rank `dst`, concatenated in rank order (works with gloo and nccl)."""
    world = dist.get_world_size(group)
    objs = [None] * world if dist.get_rank(group) == dst else None
    dist.gather_object(chunk, objs, dst=dst, group=group)
    if objs is None:
        return None
    return concat_chunks(objs)


# ---- bench.py --gpus N (N > 1) -------------------------------------------------------------------

def bench(args, wl, rank, world, dev, METRIC, UNIT, ClockSampler, algorithmic_bytes, parity=None):
    """Weak scaling of the bench workload: every rank holds its own shard of wl['n_reads'] reads
    of a genome that grows with the number of ranks (same coverage, same duplication), generated
    straight into HBM. value = k-mers of all ranks / max-over-ranks time.
    parity: callable(rank, world, local device) -> dict, run untimed before the measurement (bench.py
    passes tests/dist_check.run_cases: the real DistBuild against the oracle); a failed check is
    printed in the line and makes the process exit non-zero."""
    import json
    import os
    import time
    import torch
    import torch.distributed as dist
    from . import synth

    parity_res = parity(rank, world, dev.index) if parity is not None else None
    K, L = wl["K"], wl["read_len"]
    alphabet = wl.get("alphabet", "DNA")
    strong = wl.get("scaling") == "strong"
    if strong:      # the named input is split across the ranks (BASELINE cfg #3 / #4 / #5)
        n_total, G = wl["n_reads"], wl["G"]
        n_local = n_total // world
    else:           # cfg #2 per GPU, the genome grows with the number of ranks (same coverage)
        n_local = wl["n_reads"]
        n_total = n_local * world
        G = wl["G"] * world
    words = key_words(K, alphabet)
    # rounds of a super-batched build: one round's k-mers are held 4 x (a, the symmetric receive
    # buffer, the sort's ping-pong), so a round takes at most ~1.6 G 64-bit (0.8 G 128-bit) k-mers
    per_round = int(os.environ.get("MTG_MG_ROUND_KMERS", str(1_600_000_000 // words)))
    n_rounds = max(1, -(-(n_local * (L - K + 1)) // per_round))
    rows = -(-n_local // n_rounds)
    genome = synth.genome_torch(wl["seed"], G, dev) if alphabet == "DNA" else None
    slices = []
    for j in range(n_rounds):
        a0 = rank * n_local + j * rows
        a1 = min((rank + 1) * n_local, a0 + rows)
        if alphabet == "DNA":
            slices.append(synth.dna_reads_packed_torch(wl["seed"], n_total, L, G, wl["e"], device=dev,
                                                       r0=a0, r1=a1, genome=genome))
        else:
            slices.append(synth.protein_reads_packed_torch(wl["seed"], n_total, L, device=dev, r0=a0, r1=a1))
    del genome
    packed = slices[0]
    n_bytes = sum(p.numel() for p in slices)
    n_windows = n_local * (L - K + 1)
    db = DistBuild(K, wl["canonical"], wl["bpc"], alphabet=alphabet, device=dev.index)
    stream = torch.cuda.current_stream()

    def pusher(p):
        return lambda ctor: ctor.add_packed_device(p.data_ptr(), p.numel(), p.shape[0] * (L - K + 1))

    def step():
        return db.build(device_output=True, batches=[pusher(p) for p in slices])

    for _ in range(args.warmup):
        st = step()
    torch.cuda.synchronize()
    dist.barrier()
    sampler = ClockSampler(dev.index)
    if rank == 0:
        sampler.start()
    db.a2a_bytes, db.a2a_ms = 0, 0.0
    db.phase_events = []
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    dist.barrier()
    ev0.record(stream)
    stats = [step() for _ in range(args.steps)]
    ev1.record(stream)
    torch.cuda.synchronize()
    dist.barrier()
    ms = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    phases = {k: round(v / args.steps, 3) for k, v in db.phase_ms().items() if k != "start"}
    db.phase_events = None
    all_phases = [None] * world
    dist.all_gather_object(all_phases, phases)
    counters = torch.tensor([stats[-1]["n_input_kmers"], stats[-1]["n_real"], stats[-1]["n_rows"] - 1,
                             db.a2a_bytes, sum(s["n_kernel_launches"] for s in stats)],
                            dtype=torch.int64, device=dev)
    dist.all_reduce(counters)
    a2a = torch.tensor([db.a2a_ms], dtype=torch.float64, device=dev)
    dist.all_reduce(a2a, op=dist.ReduceOp.MAX)
    clocks = sampler.stop() if rank == 0 else None

    # end to end: host reads in (pinned), per-rank chunk out to host memory
    e2e = None
    if not args.no_e2e and n_rounds > 1:
        e2e = {"value": None, "unit": UNIT, "skipped": "end to end is measured on the single-round workloads"}
    elif not args.no_e2e:
        host = torch.empty((n_local, L), dtype=torch.uint8, pin_memory=True)
        host.copy_(packed[:, :L])
        torch.cuda.synchronize()
        bases = host.numpy().reshape(-1)
        offsets = np.arange(n_local + 1, dtype=np.uint64) * np.uint64(L)

        def step_host():
            db.rank.ctor.add_packed(bases, offsets)
            return db.build(device_output=False)

        chunk = step_host()
        torch.cuda.synchronize()
        dist.barrier()
        e_steps = max(1, min(args.steps, 3))
        t0 = time.perf_counter()
        for _ in range(e_steps):
            chunk = step_host()
        torch.cuda.synchronize()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        d2h = torch.tensor([chunk.stats["chunk_bytes"]], dtype=torch.int64, device=dev)
        dist.all_reduce(d2h)
        e2e = {"value": int(counters[0]) * e_steps / float(dt), "unit": UNIT,
               "h2d_bytes_per_step": int(n_total * (L + 1)), "d2h_bytes_per_step": int(d2h),
               "ms_per_step": 1e3 * float(dt) / e_steps, "steps": e_steps,
               "api": "per rank: mtg_push_reads(host) + mtg_mg_* phases + all-to-all, chunk to host"}
    db_use_symm = db.use_symm
    db.close()
    if rank != 0:
        if parity_res is not None and not parity_res.get("ok"):
            raise SystemExit(1)
        return
    ms_total = float(ms)
    N = int(counters[0])
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                            "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    a2a_gbs = (int(counters[3]) / world / args.steps) / (float(a2a) / args.steps * 1e-3) / 1e9 if float(a2a) else 0.0
    line = {
        "metric": METRIC, "value": N * args.steps / (ms_total * 1e-3), "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_total / args.steps,
        "higher_is_better": True, "scaling": "strong" if strong else "weak", "vs_baseline": None,
        "dtype": "u64" if words == 1 else "u128", "data": "synthetic",
        "config": {"workload": (wl["name"] + " split across the GPUs (strong scaling)") if strong else
                               (wl["name"] + " per GPU (weak scaling: %d reads, genome %d bp in total)"
                                % (n_total, G)),
                   "alphabet": alphabet, "rounds_per_build": n_rounds,
                   "n_reads_per_gpu": n_local, "read_len": L, "k": K,
                   "mode": "canonical" if wl["canonical"] else "basic", "count_width": wl["bpc"],
                   "input_kmers": N, "real_kmers": int(counters[1]), "rows": int(counters[2]) + 1,
                   "partitioning": "key range (top 8 bits of the k-mer word); all-to-all of canonical k-mers, "
                                   "of both strands of the distinct k-mers and of target nodes (%s); "
                                   "all-gather of source nodes"
                                   % ("over NVLink into peer-mapped symmetric memory: k-mers and targets by the "
                                      "stores of their scatter kernels (mtg_mg_route / mtg_mg_targets_route), "
                                      "strands by peer copies" if db_use_symm else "NCCL all_to_all_single"),
                   "l2_policy": "inputs larger than L2 (%.2f GB of keys per GPU and step)"
                                % (n_windows * 8 * words / 1e9)},
        "clocks": clocks, "e2e": e2e, "gpu_launches": int(counters[4]),
        "roofline": {"bound": "hbm", "kernel": "whole path per GPU (SURVEY 8d byte model)",
                     "achieved": None, "peak": peak, "unit": "GB/s", "frac": None, "traffic": None,
                     "alltoall": {"bytes_sent_per_gpu_per_step": int(counters[3]) / world / args.steps,
                                  "ms_per_step": float(a2a) / args.steps, "achieved_GBps": a2a_gbs,
                                  "peak_GBps": 900.0, "frac": a2a_gbs / 900.0},
                     "phases_ms_per_rank": all_phases},
        "cpu_baseline": None,
        "parity_check": parity_res,
    }
    fd = os.environ.pop("MTG_BENCH_STDOUT_FD", None)
    if fd is not None:   # bench.py parked the real stdout while NCCL could write to fd 1
        sys.stdout.flush()
        os.dup2(int(fd), 1)
    print(json.dumps(line), flush=True)
    if parity_res is not None and not parity_res.get("ok"):
        raise SystemExit("bench.py: multi-GPU parity check FAILED: %s" % parity_res.get("mismatches"))
