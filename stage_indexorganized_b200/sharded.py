"""Range-partitioned table across the GPUs of one box: one process per GPU, torch.distributed for plumbing.

Partitioning: shard g owns a contiguous range of the table in REFERENCE key order
(include/vstore/b_tree.h:99-134 of the reference); ``splitters[g]`` = largest key of shard g (g < G-1).
  * scans run shard-local on the shard that owns the start key (every shard also holds a halo of the next
    shard's first rows, so a scan that starts in the shard completes there bit-exactly) — no data-path
    collective, only the requests are routed;
  * probes: owner = #splitters < key (``stage_gpu_route``), the batch is bucketed by owner, keys travel with
    one all-to-all, every shard probes locally (probe_kernel + gather_rows_kernel), results travel back with
    a second all-to-all and are restored to request order.

The exchange logic is backend-agnostic (NCCL on GPUs, gloo in the CPU tests); the local lookup is injected.
"""
import numpy as np
import torch
import torch.distributed as dist


def bucket_by_owner(owner, world):
    """Stable bucketing: returns (perm, counts) with owner[perm] sorted ascending."""
    owner = owner.to(torch.int64)
    perm = torch.argsort(owner, stable=True)
    counts = torch.bincount(owner, minlength=world)[:world]
    return perm, counts


def exchange_counts(counts, group=None):
    """send counts [world] -> recv counts [world] (recv[j] = what rank j sends to me)."""
    recv = torch.empty_like(counts)
    dist.all_to_all_single(recv, counts, group=group)
    return recv


def all_to_all_rows(rows, send_counts, recv_counts, group=None):
    """rows: [n_send, ...] grouped by destination rank. Returns [n_recv, ...] grouped by source rank."""
    n_recv = int(recv_counts.sum().item())
    out = torch.empty((n_recv,) + tuple(rows.shape[1:]), dtype=rows.dtype, device=rows.device)
    dist.all_to_all_single(out, rows.contiguous(), output_split_sizes=[int(x) for x in recv_counts.tolist()],
                           input_split_sizes=[int(x) for x in send_counts.tolist()], group=group)
    return out


def routed_probe(keys, owner, local_probe, payload_size, group=None):
    """Distributed batched point probe.

    keys  : [n, K] uint8 tensor (this rank's requests), owner: [n] shard index per key.
    local_probe(keys_local [m, K]) -> (found u8 [m], ts i32 [m], payload u8 [m, P]) on the same device.
    Returns (found [n], ts [n], payload [n, P]) in request order.
    """
    world = dist.get_world_size(group)
    perm, send_counts = bucket_by_owner(owner, world)
    recv_counts = exchange_counts(send_counts, group)
    keys_sorted = keys.index_select(0, perm)
    keys_in = all_to_all_rows(keys_sorted, send_counts, recv_counts, group)            # requests I own
    found_l, ts_l, pay_l = local_probe(keys_in)
    # results travel back along the reverse routes
    found_b = all_to_all_rows(found_l, recv_counts, send_counts, group)
    ts_b = all_to_all_rows(ts_l, recv_counts, send_counts, group)
    pay_b = all_to_all_rows(pay_l, recv_counts, send_counts, group)
    n = keys.shape[0]
    found = torch.empty(n, dtype=found_b.dtype, device=keys.device)
    ts = torch.empty(n, dtype=ts_b.dtype, device=keys.device)
    pay = torch.empty((n, payload_size), dtype=torch.uint8, device=keys.device)
    found.index_copy_(0, perm, found_b)
    ts.index_copy_(0, perm, ts_b)
    pay.index_copy_(0, perm, pay_b)
    return found, ts, pay


def routed_scan_requests(start_keys, sizes, owner, group=None):
    """Route scan requests to the shard owning each start key. Returns (keys_in, sizes_in, perm, send_counts,
    recv_counts): the requests this rank must run, plus what is needed to send row counts back."""
    world = dist.get_world_size(group)
    perm, send_counts = bucket_by_owner(owner, world)
    recv_counts = exchange_counts(send_counts, group)
    keys_in = all_to_all_rows(start_keys.index_select(0, perm), send_counts, recv_counts, group)
    sizes_in = all_to_all_rows(sizes.index_select(0, perm), send_counts, recv_counts, group)
    return keys_in, sizes_in, perm, send_counts, recv_counts


def gather_scan_counts(row_counts_local, perm, send_counts, recv_counts, group=None):
    """The 'gather' that follows shard-local scans: per-scan row counts back to the requesters, request order."""
    back = all_to_all_rows(row_counts_local, recv_counts, send_counts, group)
    out = torch.empty_like(back)
    out.index_copy_(0, perm, back)
    return out


def shard_bounds(n_rows_total, world):
    """Equal-size range shards over ranks of the key order: [first_rank, end_rank) per shard."""
    per = (n_rows_total + world - 1) // world
    return [(min(g * per, n_rows_total), min((g + 1) * per, n_rows_total)) for g in range(world)]


def ycsb_splitters(n_rows_total, world):
    """Splitter keys (uint8 [world-1, 4]) of equal-size shards of the YCSB int table 0..N-1."""
    from .ycsb import rank_to_key
    b = shard_bounds(n_rows_total, world)
    last = np.array([b[g][1] - 1 for g in range(world - 1)], np.int64)
    return rank_to_key(n_rows_total, last).astype(np.uint32).view(np.uint8).reshape(-1, 4)


class PeerProbe:
    """Routed probes with the exchange fused into the kernels (csrc/exchange.cuh): keys are stored straight into the
    owners' inboxes and payload rows straight into the requesters' result slabs over NVLink peer memory; results
    arrive in request order. torch.distributed only carries the control plane (the one-time all-gather of the CUDA
    IPC handles, and — without device_sync — two stream-ordered barriers per batch)."""

    def __init__(self, shard, splitters, capacity, device_sync=True, group=None, max_scan_size=0):
        import stage_indexorganized_b200 as sg
        self.sg, self.shard, self.group = sg, shard, group
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.splitters = splitters
        self.flags = sg.XCHG_DEVICE_SYNC if device_sync else 0
        self.x = sg.ProbeExchange(self.world, self.rank, capacity, shard.schema.key_size, shard.schema.payload_size,
                                  max_scan_size=max_scan_size)
        handles = [None] * self.world
        dist.all_gather_object(handles, self.x.ipc_handle(), group=group)
        if self.world > 1:
            self.x.attach_ipc(handles)
        self._tok = torch.zeros(1, dtype=torch.int32, device=torch.device("cuda", torch.cuda.current_device()))
        dist.barrier(group)          # every arena is mapped everywhere before the first post

    def _barrier(self):
        dist.all_reduce(self._tok, group=self.group)      # stream-ordered; no host synchronisation

    def probe(self, keys_dev, n, read_ts):
        """keys_dev: [n, K] uint8 CUDA tensor. Returns (found, ts, payload) views of this rank's result slab."""
        st = torch.cuda.current_stream().cuda_stream
        self.x.post(self.splitters, keys_dev, n, self.flags, st)
        if not self.flags:
            self._barrier()
        self.x.serve(self.shard, read_ts, self.flags, st)
        if self.flags:
            self.x.wait(st)
        else:
            self._barrier()
        return self.x.results_torch(n)

    def measure_link_ceiling(self, reps=3):
        """GB/s per direction per GPU that the copy engines reach on THIS box for the exchange's traffic pattern: every rank
        copies an equal slice of its slab into every other rank's slab, all ranks at once (best of `reps`, max over
        ranks). The slab contents are garbage afterwards."""
        if self.world == 1:
            return None
        _, _, _, stride = self.x.result_ptrs()
        total = self.x.capacity * stride
        per = (total // self.world) // 4096 * 4096
        stream = torch.cuda.current_stream()
        best = None
        for _ in range(reps + 1):
            torch.cuda.synchronize()
            dist.barrier(self.group)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for d in range(1, self.world):
                self.x.peer_copy((self.rank + d) % self.world, self.rank * per, per, stream.cuda_stream)
            e1.record(stream)
            torch.cuda.synchronize()
            ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=torch.device("cuda", torch.cuda.current_device()))
            dist.all_reduce(ms, op=dist.ReduceOp.MAX, group=self.group)
            best = float(ms.item()) if best is None else min(best, float(ms.item()))
        dist.barrier(self.group)
        return per * (self.world - 1) / (best * 1e-3) / 1e9

    def probe_column(self, keys_dev, n, read_ts, col_off, col_len):
        """Routed probes with column push-down: only payload bytes [col_off, col_off + col_len) of every found row cross
        NVLink. Returns (found, ts, col[n, col_len]) views of this rank's result slab."""
        st = torch.cuda.current_stream().cuda_stream
        self.x.post(self.splitters, keys_dev, n, self.flags, st)
        if not self.flags:
            self._barrier()
        self.x.serve_project(self.shard, read_ts, col_off, col_len, self.flags, st)
        if self.flags:
            self.x.wait(st)
        else:
            self._barrier()
        return self.x.project_results_torch(n, col_len)

    def probe_aggregate_i32(self, keys_dev, n, read_ts, col_off, threshold):
        """Routed probes reduced on the owners: AggI32(found, below, min_value, max_value) over this rank's batch; one
        24-byte record per (owner, requester) pair crosses NVLink."""
        st = torch.cuda.current_stream().cuda_stream
        self.x.post(self.splitters, keys_dev, n, self.flags, st)
        if not self.flags:
            self._barrier()
        self.x.serve_aggregate_i32(self.shard, read_ts, col_off, threshold, self.flags, st)
        if self.flags:
            self.x.wait(st)
        else:
            self._barrier()
        return self.x.aggregate_result(st)

    def close(self):
        torch.cuda.synchronize()
        dist.barrier(self.group)     # nobody unmaps an arena a peer may still be writing
        self.x.close()


class PeerScan(PeerProbe):
    """Routed range scans over the same arenas: each scan runs on the owner of its start key (whose shard carries a
    halo of its right neighbour's first rows) and its rows are written into this rank's slab in request order."""

    def __init__(self, shard, splitters, max_scans, max_scan_size, device_sync=True, group=None):
        super().__init__(shard, splitters, max_scans, device_sync, group, max_scan_size=max(2, max_scan_size))

    def scan(self, start_keys_dev, scan_sizes_dev, n, read_ts, apply_visibility=True):
        """Returns (row_counts[n], keys[n,R,K], ts[n,R], payload[n,R,P]) views of this rank's result slab."""
        st = torch.cuda.current_stream().cuda_stream
        self.x.post_scans(self.splitters, start_keys_dev, scan_sizes_dev, n, self.flags, st)
        if not self.flags:
            self._barrier()
        self.x.serve_scans(self.shard, read_ts, apply_visibility, self.flags, st)
        if self.flags:
            self.x.wait(st)
        else:
            self._barrier()
        return self.x.scan_results_torch(n)
