"""Range-partitioned BOSS chunk build across the GPUs of one box.

One process per GPU (torch.distributed, NCCL over NVLink/NVSwitch for the plumbing). The path is
partitioned the way the reference's own chunked build partitions it -- by the leading characters
of the sort key (boss_chunk_construct.cpp:366-397, :462-525, :754-792, :917-931):

  1. every rank extracts the k-mers of its read shard and groups them by routing bucket
     (top part_bits bits of the KMerBOSS word)                         mtg_mg_extract
  2. all-to-all #1: every bucket goes to the rank that owns its range   NCCL all_to_all_single
  3. owners sort + fold their k-mers (+ counts)                          mtg_mg_collect
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


def concat_chunks(chunks):
    """Host concatenation, exactly BOSS::Chunk::extend in rank (= key range) order."""
    full = chunks[0]
    for c in chunks[1:]:
        full.extend(c)
    return full


# ---- device buffers ---------------------------------------------------------------------------

class _CudaPtr:
    """Zero-copy torch view of library-owned device memory (__cuda_array_interface__)."""

    def __init__(self, ptr, n, words):
        shape = (n, words) if words > 1 else (n,)
        self.__cuda_array_interface__ = {"shape": shape, "typestr": "<i8", "data": (ptr, False),
                                         "version": 2, "strides": None}


def as_tensor(ptr, n, words, device):
    import torch
    if n == 0:
        return torch.empty((0, words) if words > 1 else (0,), dtype=torch.int64, device=device)
    return torch.as_tensor(_CudaPtr(ptr, n, words), device=device)


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
        self.ctor = BOSSChunkConstructor.initialize(K - 1, canonical, bpc, alphabet=alphabet, device=device)
        self.nb = 1 << self.ctor.mg_part_bits()

    def extract(self):
        counts, ptr, n = self.ctor.mg_extract()
        self.n_extracted = n
        return counts, as_tensor(ptr, n, self.words, self.device)

    def collect(self, keys):
        self._keep = keys   # the library copies; keep the tensor alive until the call returns
        n = keys.shape[0]
        self.n_real = self.ctor.mg_collect(keys.data_ptr() if n else 0, n)
        self._keep = None
        return self.n_real

    def targets(self):
        counts, ptr, n = self.ctor.mg_targets()
        return counts, as_tensor(ptr, n, self.words, self.device)

    def dummies(self, targets_in):
        n = targets_in.shape[0]
        ptr, n_src, n_sink = self.ctor.mg_dummies(targets_in.data_ptr() if n else 0, n)
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

    def __init__(self, world, K, canonical=False, bits_per_count=0, alphabet="DNA", device=0):
        self.world = world
        self.ranks = [RangedRank(K, canonical, bits_per_count, alphabet, device) for _ in range(world)]

    def add_sequences(self, rank, sequences):
        self.ranks[rank].ctor.add_sequences(sequences)

    def build_chunks(self):
        torch = self.ranks[0].torch
        g = self.world
        ex = [r.extract() for r in self.ranks]
        hist = np.sum([c for c, _ in ex], axis=0)
        self.ranges = plan_ranges(hist, g)
        offs = [_bucket_offsets(c, self.ranges) for c, _ in ex]
        routed = [torch.cat([ex[s][1][offs[s][d][0]:offs[s][d][1]] for s in range(g)]) for d in range(g)]
        for d, r in enumerate(self.ranks):
            r.collect(routed[d])
        del routed
        tg = [r.targets() for r in self.ranks]
        toffs = [_bucket_offsets(c, self.ranges) for c, _ in tg]
        recv = [torch.cat([tg[s][1][toffs[s][d][0]:toffs[s][d][1]] for s in range(g)]) for d in range(g)]
        src = [r.dummies(recv[d]).clone() for d, r in enumerate(self.ranks)]
        all_src = torch.cat(src)
        chunks = [r.build(all_src, self.ranges[d][0], self.ranges[d][1], d == 0)
                  for d, r in enumerate(self.ranks)]
        return chunks

    def build_chunk(self):
        return concat_chunks(self.build_chunks())

    def close(self):
        for r in self.ranks:
            r.close()


# ---- one process per GPU ------------------------------------------------------------------------

class DistBuild:
    """The partitioned build of THIS rank inside a torch.distributed (NCCL) group."""

    def __init__(self, K, canonical=False, bits_per_count=0, alphabet="DNA", device=0, group=None):
        import torch
        import torch.distributed as dist
        self.torch, self.dist, self.group = torch, dist, group
        self.world = dist.get_world_size(group)
        self.rank_id = dist.get_rank(group)
        self.rank = RangedRank(K, canonical, bits_per_count, alphabet, device)
        self.a2a_bytes = 0
        self.a2a_ms = 0.0

    def _exchange(self, send, send_sizes):
        """all_to_all of a bucket-grouped tensor; send_sizes[d] elements go to rank d."""
        torch, dist = self.torch, self.dist
        dev = self.rank.device
        sizes = torch.tensor(send_sizes, dtype=torch.int64, device=dev)
        all_sizes = torch.empty((self.world, self.world), dtype=torch.int64, device=dev)
        dist.all_gather_into_tensor(all_sizes, sizes, group=self.group)
        recv_sizes = all_sizes[:, self.rank_id].tolist()
        shape = (int(sum(recv_sizes)),) + tuple(send.shape[1:])
        recv = torch.empty(shape, dtype=torch.int64, device=dev)
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        dist.all_to_all_single(recv, send.contiguous(), recv_sizes, list(send_sizes), group=self.group)
        ev1.record()
        ev1.synchronize()
        self.a2a_ms += ev0.elapsed_time(ev1)
        own = send_sizes[self.rank_id]
        self.a2a_bytes += (int(sum(send_sizes)) - own) * 8 * self.rank.words
        return recv

    def build(self, device_output=False):
        torch, dist = self.torch, self.dist
        dev = self.rank.device
        counts, keys = self.rank.extract()
        hist = torch.from_numpy(counts.astype(np.int64)).to(dev)
        dist.all_reduce(hist, group=self.group)
        self.ranges = plan_ranges(hist.cpu().numpy(), self.world)
        self.rank.collect(self._exchange(keys, split_sizes(counts, self.ranges)))
        tcounts, tg = self.rank.targets()
        src = self.rank.dummies(self._exchange(tg, split_sizes(tcounts, self.ranges)))
        n_src = torch.tensor([src.shape[0]], dtype=torch.int64, device=dev)
        all_n = torch.empty(self.world, dtype=torch.int64, device=dev)
        dist.all_gather_into_tensor(all_n, n_src, group=self.group)
        all_n = all_n.tolist()
        parts = [torch.empty((n,) + tuple(src.shape[1:]), dtype=torch.int64, device=dev) for n in all_n]
        self._all_gather_ragged(parts, src, all_n)
        all_src = torch.cat(parts)
        lo, hi = self.ranges[self.rank_id]
        return self.rank.build(all_src, lo, hi, self.rank_id == 0, device_output)

    def _all_gather_ragged(self, parts, src, all_n):
        dist = self.dist
        for r in range(self.world):   # broadcast per rank: the lists are tiny
            buf = src.contiguous() if r == self.rank_id else parts[r]
            if all_n[r]:
                dist.broadcast(buf, src=dist.get_global_rank(self.group, r) if self.group else r,
                               group=self.group)
            parts[r] = buf

    def close(self):
        self.rank.close()


def gather_chunks(chunk, dist, group=None, dst=0):
    """Per-rank chunks -> rank `dst`, concatenated in rank order (works with gloo and nccl)."""
    world = dist.get_world_size(group)
    objs = [None] * world if dist.get_rank(group) == dst else None
    dist.gather_object(chunk, objs, dst=dst, group=group)
    if objs is None:
        return None
    return concat_chunks(objs)
