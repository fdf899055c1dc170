"""Data-parallel plumbing for the quantiser: one process per GPU, replicated codebook,
row-sharded latents, ONE collective per step -- all_reduce(SUM) over the contiguous
[code_sum (K*D) | code_cnt (K)] fp32 buffer -- after which every rank applies the identical EMA update.
The reference has no distributed code (single GPU, frl/config/frl_training_v1.yaml:26-27); this is the
shard point north_star prescribes.  Backend-agnostic (NCCL on the GPUs, gloo in the CPU tests)."""
from __future__ import annotations

import os
from typing import Optional, Tuple

import torch
import torch.distributed as dist


def shard_bounds(n_rows: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous, balanced row shard [a, b) of rank `rank` (first n_rows % world ranks get one more)."""
    base, rem = divmod(int(n_rows), int(world))
    a = rank * base + min(rank, rem)
    return a, a + base + (1 if rank < rem else 0)


def pin_to_gpu_numa(local_rank: int):
    """Bind this process (and the pinned host memory it allocates from now on: first touch) to the CPUs of the NUMA node
    its GPU hangs off, so that eight ranks staging batches through pinned memory do not all pull from node 0.  Reads
    the PCI bus id from CUDA and the node from sysfs; returns a short description, or None when nothing could be done
    (no sysfs entry, node -1, a container without the cpulist).  Call before the first pinned allocation."""
    try:
        import torch
        bus = torch.cuda.get_device_properties(local_rank).pci_bus_id
        dom = torch.cuda.get_device_properties(local_rank).pci_domain_id
        dev = torch.cuda.get_device_properties(local_rank).pci_device_id
        path = f"/sys/bus/pci/devices/{dom:04x}:{bus:02x}:{dev:02x}.0/numa_node"
        with open(path) as fh:
            node = int(fh.read().strip())
        if node < 0:
            return None
        with open(f"/sys/devices/system/node/node{node}/cpulist") as fh:
            spec = fh.read().strip()
        cpus = set()
        for part in spec.split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return f"gpu {local_rank} -> numa node {node} ({len(cpus)} cpus)"
    except Exception:
        return None


def dist_active(group=None) -> bool:
    return dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1


def allreduce_code_stats(stats: torch.Tensor, group=None) -> torch.Tensor:
    """In-place SUM over ranks of the packed statistics buffer; a no-op for a single process."""
    if dist_active(group):
        dist.all_reduce(stats, op=dist.ReduceOp.SUM, group=group)
    return stats


def codebook_checksum(codebook: torch.Tensor) -> torch.Tensor:
    """Order-independent fingerprint used to assert that replicas have not drifted apart."""
    bits = codebook.detach().contiguous().view(torch.int32).to(torch.int64)
    return torch.stack([bits.sum(), (bits * bits % 1000003).sum()])


def assert_replicas_identical(codebook: torch.Tensor, group=None) -> None:
    if not dist_active(group):
        return
    mine = codebook_checksum(codebook)
    world = dist.get_world_size(group)
    allc = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(allc, mine, group=group)
    for r, c in enumerate(allc):
        if not torch.equal(c, allc[0]):
            raise RuntimeError(f"codebook replica of rank {r} differs from rank 0")


class _RawCudaBuffer:
    """Exposes raw device memory to torch through __cuda_array_interface__ (no copy, no ownership)."""

    def __init__(self, ptr: int, numel: int, typestr: str):
        self.__cuda_array_interface__ = {"shape": (numel,), "typestr": typestr, "data": (ptr, False), "version": 2}


class SymmExchange:
    """Statistics exchange buffers on torch symmetric memory (plumbing only: allocation + rendezvous): same layout and
    interface as PeerExchange -- [stats 0 | stats 1 | flags] per rank, peer pointers for the flags -- plus, where the
    node has NVLS, the MULTICAST address of the allocation: the EMA kernels then read every statistic with one
    multimem.ld_reduce (sum over ranks formed inside the NVSwitch, one NVLink read per element instead of W).
    reduced_numel > 0 adds a [reduced_numel] float region behind the flags for the two-phase form (include/vq_b200.h,
    vq_ema_update_dp: mc_reduced / reduced_local), used for codebooks of TWO_PHASE_MIN_BYTES and more."""

    TWO_PHASE_MIN_BYTES = 2 << 20        # K*D*4; below it the one-phase form (one kernel, one flag round less) wins

    def __init__(self, numel: int, device: torch.device, group=None, timeout_s: Optional[float] = None,
                 reduced_numel: int = 0):
        import torch.distributed._symmetric_memory as symm
        from . import _lib
        lib = _lib.load()
        if timeout_s is None:
            timeout_s = float(os.environ.get("VQ_PEER_TIMEOUT_S", "1800"))
        lib.vq_set_peer_timeout_ms(int(timeout_s * 1000))
        self.group, self.numel, self.device = group, int(numel), device
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        if self.world > _lib.MAX_PEERS:
            raise RuntimeError(f"SymmExchange supports at most {_lib.MAX_PEERS} ranks")
        stats_bytes = (self.numel * 4 + 255) // 256 * 256
        self._stats_bytes = stats_bytes
        self._red_numel = int(reduced_numel)
        total_f = (2 * stats_bytes + 256) // 4 + self._red_numel
        g = group if group is not None else dist.group.WORLD
        self._t = symm.empty(total_f, dtype=torch.float32, device=device)
        self._h = symm.rendezvous(self._t, g.group_name)
        self._t.zero_()
        self._peer_base = [int(p) for p in self._h.buffer_ptrs]
        mc = getattr(self._h, "multicast_ptr", 0) or 0
        self._mc_base = int(mc) if os.environ.get("VQ_PEER_MULTICAST", "1") != "0" else 0
        self.stats = [self._t[i * stats_bytes // 4: i * stats_bytes // 4 + self.numel] for i in range(2)]
        self.epoch = 0
        tp = os.environ.get("VQ_PEER_TWO_PHASE", "auto")
        self._two_phase = bool(self._mc_base) and self._red_numel > 0 and tp != "0" and \
            (tp == "1" or self._red_numel * 4 >= self.TWO_PHASE_MIN_BYTES)
        self.kind = ("symm+multicast, two-phase" if self._two_phase else "symm+multicast") if self._mc_base else "symm"
        torch.cuda.synchronize(device)
        dist.barrier(group=group)            # every allocation is zeroed and mapped before the first signal

    def next_buffer(self) -> torch.Tensor:
        self.epoch += 1
        return self.stats[self.epoch & 1]

    def peer_stats(self):
        return [b + (self.epoch & 1) * self._stats_bytes for b in self._peer_base]

    def peer_flags(self):
        return [b + 2 * self._stats_bytes for b in self._peer_base]

    def mc_stats(self):
        return self._mc_base + (self.epoch & 1) * self._stats_bytes if self._mc_base else None

    def reduced(self):
        """(multicast address, this rank's address) of the reduced-sums region, or (None, None): one-phase exchange."""
        if not self._two_phase:
            return None, None
        off = 2 * self._stats_bytes + 256
        return self._mc_base + off, self._t.data_ptr() + off

    def close(self) -> None:
        if self._t is None:
            return
        torch.cuda.synchronize()
        if dist.is_initialized():
            dist.barrier(group=self.group)
        self._t, self._h, self.stats = None, None, []


def make_exchange(numel: int, device: torch.device, group=None, timeout_s: Optional[float] = None,
                  reduced_numel: int = 0):
    """Symmetric-memory exchange (with the in-switch reduction where the node has NVLS) when torch provides it, else the
    CUDA-IPC one.  VQ_PEER_SYMM=0 forces the IPC path.  Every rank takes the same branch (the choice is all-reduced)."""
    ok = os.environ.get("VQ_PEER_SYMM", "1") != "0"
    ex = None
    if ok:
        try:
            ex = SymmExchange(numel, device, group, timeout_s, reduced_numel)
        except Exception as exc:          # no symmetric memory on this build / node: the IPC path does the same job
            import warnings
            warnings.warn(f"vqvae_b200: symmetric-memory exchange unavailable ({type(exc).__name__}: {exc}); using CUDA IPC")
            ex = None
    flag = torch.tensor([1 if ex is not None else 0], device=device)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
    if int(flag) == 1:
        return ex
    if ex is not None:
        ex.close()
    return PeerExchange(numel, device, group, timeout_s)


class PeerExchange:
    """Peer-memory statistics exchange for the fused a-6 + a-7 update (one process per GPU, one node).

    Every rank owns ONE cudaMalloc'd allocation [stats 0 | stats 1 | flags] exported through CUDA IPC; the handles
    travel once over the process group (all_gather_object), each rank maps its peers' allocations, and from then
    on a step needs no collective call: the pass kernel writes the local statistics buffer of the step's parity,
    vq_ema_update_dp signals / waits on the flags and reads all ranks' buffers over NVLink.
    Two statistics buffers alternate because a slower rank may still be reading step t while a faster one already
    writes step t+1 (the flag exchange of step t+1 orders everything older).
    """

    def __init__(self, numel: int, device: torch.device, group=None, timeout_s: Optional[float] = None):
        import ctypes as C
        from . import _lib
        lib = _lib.load()
        # the cross-rank wait inside the kernels is a rendezvous of processes: bound it like a collective (minutes), not
        # like a kernel-internal barrier.  VQ_PEER_TIMEOUT_S / timeout_s; default 30 min; 0 = wait for ever.
        if timeout_s is None:
            timeout_s = float(os.environ.get("VQ_PEER_TIMEOUT_S", "1800"))
        lib.vq_set_peer_timeout_ms(int(timeout_s * 1000))
        self.group, self.numel, self.device = group, int(numel), device
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        if self.world > _lib.MAX_PEERS:
            raise RuntimeError(f"PeerExchange supports at most {_lib.MAX_PEERS} ranks")
        stats_bytes = (self.numel * 4 + 255) // 256 * 256
        self._stats_bytes = stats_bytes
        total = 2 * stats_bytes + 256
        base, handle = C.c_void_p(), C.create_string_buffer(_lib.IPC_HANDLE_BYTES)
        _lib.check(lib.vq_ipc_alloc(total, C.byref(base), handle), "vq_ipc_alloc")
        self._base, self._lib, self._opened = int(base.value), lib, []
        handles = [None] * self.world
        dist.all_gather_object(handles, (os.getpid(), bytes(handle.raw)), group=group)
        self._peer_base = []
        for r, (pid, h) in enumerate(handles):
            if r == self.rank:
                self._peer_base.append(self._base)
                continue
            if pid == os.getpid():
                raise RuntimeError("PeerExchange needs one process per GPU")
            p = C.c_void_p()
            _lib.check(lib.vq_ipc_open(C.create_string_buffer(h, _lib.IPC_HANDLE_BYTES), C.byref(p)), "vq_ipc_open")
            self._opened.append(int(p.value))
            self._peer_base.append(int(p.value))
        self.stats = [torch.as_tensor(_RawCudaBuffer(self._base + i * stats_bytes, self.numel, "<f4"), device=device)
                      for i in range(2)]
        self.epoch = 0
        self.kind = "ipc"
        dist.barrier(group=group)            # every allocation is zeroed and mapped before the first signal

    def mc_stats(self):
        return None

    def reduced(self):
        return None, None

    def next_buffer(self) -> torch.Tensor:
        """Statistics buffer the NEXT step writes (call once per step, before the pass kernel)."""
        self.epoch += 1
        return self.stats[self.epoch & 1]

    def peer_stats(self):
        return [b + (self.epoch & 1) * self._stats_bytes for b in self._peer_base]

    def peer_flags(self):
        return [b + 2 * self._stats_bytes for b in self._peer_base]

    def close(self) -> None:
        if self._lib is None:
            return
        torch.cuda.synchronize()
        if dist.is_initialized():
            dist.barrier(group=self.group)   # nobody still reads a buffer that is about to be freed
        for p in self._opened:
            self._lib.vq_ipc_close(p)
        self._lib.vq_ipc_free(self._base)
        self._lib, self.stats = None, []

    def __del__(self):
        try:
            if self._lib is not None:
                for p in self._opened:
                    self._lib.vq_ipc_close(p)
                self._lib.vq_ipc_free(self._base)
        except Exception:
            pass
