"""Multi-GPU host layer (one process per GPU, torch.distributed for the plumbing).

Two partitionings of the hot path (the reference itself is single-process, its only
parallelism is hybrid_loop's std::thread fan-out, /root/reference/include/gaussianblur/helpers.hpp:134-172):

* frames  -> ranks, contiguous blocks, NO collective (video batches; every rank holds the same plan);
* row bands of one image -> ranks: row pass local, ONE all-to-all of the fp32 intermediate so that every
  rank then owns a band of pixel columns for all rows, column pass local. The all-to-all is NCCL
  (`dist.all_to_all`) on GPUs; on backends without it (gloo, used by the CPU tests of this plumbing) the
  same exchange is done with point-to-point sends.

The compute is injected (`rowpass` / `colpass` callables): the product binds them to the CUDA plans
(`BandShardedBlur.cuda`), the CPU tests bind them to the test oracle to check partitioning, packing and
exchange without a GPU. There is no CPU compute path in this module.
"""
from __future__ import annotations

import os
from typing import Callable, List, Sequence, Tuple

import torch
import torch.distributed as dist


def shard_range(n: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous block [start, stop) of n items owned by `rank`; the first n % world ranks get one extra."""
    base, rem = divmod(n, world)
    start = rank * base + min(rank, rem)
    return start, start + base + (1 if rank < rem else 0)


def shard_ranges(n: int, world: int) -> List[Tuple[int, int]]:
    return [shard_range(n, world, r) for r in range(world)]


def exchange(send: Sequence[torch.Tensor], recv: Sequence[torch.Tensor], group=None) -> None:
    """all-to-all of per-peer tensors (send[h] goes to rank h, recv[g] comes from rank g)."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        recv[0].copy_(send[0])
        return
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    if dist.get_backend(group) == "nccl":
        dist.all_to_all(list(recv), list(send), group=group)
        return
    # gloo & co.: pairwise exchange, lower rank sends first (deadlock-free for blocking send/recv)
    recv[rank].copy_(send[rank])
    for peer in range(world):
        if peer == rank:
            continue
        src = dist.get_global_rank(group, peer) if group is not None else peer
        if rank < peer:
            dist.send(send[peer].contiguous(), src, group=group)
            dist.recv(recv[peer], src, group=group)
        else:
            dist.recv(recv[peer], src, group=group)
            dist.send(send[peer].contiguous(), src, group=group)


def panel_exchange_ok(cols: int, cp: int, world: int) -> bool:
    """Can the fused / staged exchange (row pass stores into the owners' panel buffers) be used? It splits the float
    columns evenly over the owners, so the pixel-column bands of shard_range must be equal and start on panel (32-float)
    boundaries: cols = 64, world = 3, RGB passes the first condition alone and would shift columns."""
    return (cols * cp) % (32 * world) == 0 and cols % world == 0 and ((cols // world) * cp) % 32 == 0 and world <= 8


class BandShardedBlur:
    """One (rows x cols x channels) image sharded by row bands over the ranks of `group`.

    rowpass(band_u8[band_rows, cols, C]) -> fp32 [band_rows, cols * Cp]          (local rows, all columns)
    colpass(inter[rows, band_cols * Cp], alpha_src) -> u8 [rows, band_cols, C]   (all rows, local columns)
    run() returns the rank's band of pixel COLUMNS of the blurred image (all rows); `gather_rows=True`
    adds a second (4x smaller, u8) all-to-all that returns to row-band ownership.
    """

    def __init__(self, rows: int, cols: int, channels: int, processed_channels: int,
                 rowpass: Callable[[torch.Tensor], torch.Tensor],
                 colpass: Callable[[torch.Tensor, torch.Tensor], torch.Tensor], group=None):
        self.rows, self.cols, self.C, self.Cp = rows, cols, channels, processed_channels
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.row_ranges = shard_ranges(rows, self.world)
        self.col_ranges = shard_ranges(cols, self.world)
        self.rowpass, self.colpass = rowpass, colpass
        self.rowpass_split = None  # optional: band -> [world, band_rows, cols/world, Cp] written in place
        self.profile = False       # fused path: record CUDA events around fence / row pass / fence / column pass
        self.marks = []
        self.fused_run = None      # optional: band -> column band, exchange fused into the row pass (peer stores)
        self.exchanged_bytes = 0  # bytes this rank sent to OTHER ranks in the last run

    @property
    def my_rows(self) -> Tuple[int, int]:
        return self.row_ranges[self.rank]

    @property
    def my_cols(self) -> Tuple[int, int]:
        return self.col_ranges[self.rank]

    def run(self, band: torch.Tensor, gather_rows: bool = False) -> torch.Tensor:
        if self.fused_run is not None:
            out_cols = self.fused_run(band)
            return self._gather_rows(out_cols) if gather_rows else out_cols
        r0, r1 = self.my_rows
        c0, c1 = self.my_cols
        Cp, C = self.Cp, self.C
        assert band.shape[0] == r1 - r0 and band.shape[1] == self.cols
        if self.rowpass_split is not None and self.cols % self.world == 0:
            # the row pass writes the per-destination blocks itself: no packing copy
            send = list(self.rowpass_split(band).unbind(0))
            inter = send[0]
        else:
            inter = self.rowpass(band).view(r1 - r0, self.cols, Cp)
            # pack: one contiguous block per destination rank (its pixel columns of my rows)
            send = [inter[:, a:b, :].contiguous() for a, b in self.col_ranges]
        recv_buf = torch.empty((self.rows, c1 - c0, Cp), dtype=inter.dtype, device=inter.device)
        recv = [recv_buf[a:b] for a, b in self.row_ranges]  # row blocks of a row-major buffer: contiguous
        self.exchanged_bytes = sum(t.numel() * t.element_size() for h, t in enumerate(send) if h != self.rank)
        exchange(send, recv, self.group)
        # the un-blurred alpha bytes (C == 4, Cp == 3) travel as u8 beside the intermediate
        alpha_src = None
        if Cp < C:
            a_send = [band[:, a:b, C - 1].contiguous() for a, b in self.col_ranges]
            alpha_src = torch.empty((self.rows, c1 - c0), dtype=band.dtype, device=band.device)
            exchange(a_send, [alpha_src[a:b] for a, b in self.row_ranges], self.group)
        out_cols = self.colpass(recv_buf.view(self.rows, (c1 - c0) * Cp), alpha_src)
        return self._gather_rows(out_cols) if gather_rows else out_cols

    def _gather_rows(self, out_cols: torch.Tensor) -> torch.Tensor:
        """Second, 4x smaller (u8) all-to-all: from column-band ownership back to row bands."""
        r0, r1 = self.my_rows
        C = self.C
        back_send = [out_cols[a:b].contiguous() for a, b in self.row_ranges]
        parts = [torch.empty((r1 - r0, b - a, C), dtype=out_cols.dtype, device=out_cols.device)
                 for a, b in self.col_ranges]
        exchange(back_send, parts, self.group)
        return torch.cat(parts, dim=1)

    # ---- product binding: CUDA plans through the C ABI ----
    @classmethod
    def cuda(cls, rows: int, cols: int, channels: int, sigma: float, apply_to_alpha: bool = False, device: int = 0,
             group=None, fused: bool = True, pipeline_chunks: int = 0) -> "BandShardedBlur":
        import libgaussianblur_b200 as gb

        world = dist.get_world_size(group) if dist.is_initialized() else 1
        rank = dist.get_rank(group) if dist.is_initialized() else 0
        r0, r1 = shard_range(rows, world, rank)
        c0, c1 = shard_range(cols, world, rank)
        cp = 3 if (channels == 4 and not apply_to_alpha) else channels
        p_row = gb.Plan(rows, cols, channels, sigma, device=device, band_rows=r1 - r0, band_cols=cols)
        p_col = gb.Plan(rows, cols, channels, sigma, device=device, band_rows=rows, band_cols=c1 - c0)
        dev = torch.device("cuda", device)
        wb = (c1 - c0) * cp

        def rowpass(band: torch.Tensor) -> torch.Tensor:
            inter = torch.empty((r1 - r0, cols * cp), dtype=torch.float32, device=dev)
            p_row.rowpass(band, inter, cols * cp, apply_to_alpha)
            return inter

        def colpass(inter: torch.Tensor, alpha_src) -> torch.Tensor:
            pitch = (wb + 3) // 4 * 4
            if pitch != wb:  # the kernels read 16-byte pieces: pad the pitch
                padded = torch.zeros((rows, pitch), dtype=torch.float32, device=dev)
                padded[:, :wb] = inter
                inter = padded
            out = torch.empty((rows, c1 - c0, channels), dtype=torch.uint8, device=dev)
            if alpha_src is not None:
                out[..., channels - 1] = alpha_src
            p_col.colpass(inter, pitch, out, (c1 - c0) * channels, apply_to_alpha)
            return out

        def rowpass_split(band: torch.Tensor) -> torch.Tensor:
            bc = cols // world
            out = torch.empty((world, r1 - r0, bc, cp), dtype=torch.float32, device=dev)
            p_row.rowpass_split(band, out, bc, (r1 - r0) * bc * cp, apply_to_alpha)
            return out

        self = cls(rows, cols, channels, cp, rowpass, colpass, group)
        self.rowpass_split = rowpass_split
        self.plans = (p_row, p_col)
        self._opened = []
        if (not fused) and pipeline_chunks > 1 and world > 1 and cols % world == 0 and (r1 - r0) >= pipeline_chunks:
            # NCCL all-to-all pipelined under the row pass: the band's rows go in K chunks; chunk k's blocks
            # travel on a side stream while chunk k+1 is being computed.
            K = pipeline_chunks
            bc = cols // world
            chunk_plans = []
            for k in range(K):
                a, b = shard_range(r1 - r0, K, k)
                chunk_plans.append(gb.Plan(rows, cols, channels, sigma, device=device, band_rows=b - a, band_cols=cols))
            side = torch.cuda.Stream(device=dev)
            row_ranges = self.row_ranges

            def pipelined_run(band: torch.Tensor) -> torch.Tensor:
                comp = torch.cuda.current_stream()
                recv_buf = torch.empty((rows, bc, cp), dtype=torch.float32, device=dev)
                recv_buf.record_stream(side)
                for k, pk in enumerate(chunk_plans):
                    a, b = shard_range(r1 - r0, K, k)
                    send = torch.empty((world, b - a, bc, cp), dtype=torch.float32, device=dev)
                    pk.rowpass_split(band[a:b], send, bc, (b - a) * bc * cp, apply_to_alpha)
                    ev = torch.cuda.Event()
                    ev.record(comp)
                    send.record_stream(side)
                    recv = []
                    for g, (g0, g1) in enumerate(row_ranges):  # rank g cuts ITS band the same way
                        ga, gb_ = shard_range(g1 - g0, K, k)
                        recv.append(recv_buf[g0 + ga:g0 + gb_])
                    with torch.cuda.stream(side):
                        side.wait_event(ev)
                        exchange(list(send.unbind(0)), recv, group)
                comp.wait_stream(side)
                alpha_src = self._exchange_alpha(band) if cp < channels else None
                return colpass(recv_buf.view(rows, bc * cp), alpha_src)

            self.fused_run = pipelined_run
            self.plans = (p_row, p_col, *chunk_plans)
            self.exchange_kind = f"NCCL all-to-all in {K} row chunks, pipelined under the row pass"
            return self
        # the panel exchange splits the float columns evenly over the owners: the pixel-column bands must be equal and
        # start on panel boundaries (cols = 64, world = 3, RGB passes the first test alone and would shift columns)
        if fused and panel_exchange_ok(cols, cp, world):
            # Fused exchange: every rank's row pass stores its lines straight into the panel buffer of the rank
            # that owns those columns (CUDA IPC mapping of the peers' buffers, NVLink stores). Two tiny
            # all-reduces on the compute stream order the passes across ranks: nobody overwrites a panel that
            # its owner is still reading, and nobody reads a panel before every row pass has finished.
            panel_ptr, _ = p_col.panel_alloc(apply_to_alpha)
            peers = [panel_ptr]
            if world > 1:
                handles = [None] * world
                dist.all_gather_object(handles, gb.ipc_export(panel_ptr), group=group)
                peers = []
                for h, hd in enumerate(handles):
                    if h == rank:
                        peers.append(panel_ptr)
                    else:
                        q = gb.ipc_open(hd)
                        self._opened.append(q)
                        peers.append(q)
            token = torch.zeros(1, dtype=torch.float32, device=dev)

            def fence():
                if world > 1:
                    dist.all_reduce(token, group=group)

            def mark():
                if self.profile:
                    e = torch.cuda.Event(enable_timing=True)
                    e.record()
                    self.marks.append(e)

            staged = fused in ("staged", "hybrid") and world > 1
            hybrid = fused == "hybrid"  # every other chunk stores straight into the peers: SM stores and copy engines share the load
            if staged:
                # Staged exchange: the row pass stores into LOCAL buffers with each destination's panel geometry,
                # in K row chunks; copy engines push chunk k's slots into the owners' panel buffers over NVLink
                # while chunk k+1 is computed. The SMs never wait on remote stores.
                K = max(1, min(pipeline_chunks or 4, r1 - r0))
                local = [panel_ptr if h == rank else p_col.panel_alloc(apply_to_alpha)[0] for h in range(world)]
                chunk_plans = [gb.Plan(rows, cols, channels, sigma, device=device,
                                       band_rows=shard_range(r1 - r0, K, k)[1] - shard_range(r1 - r0, K, k)[0],
                                       band_cols=cols) for k in range(K)]
                self.plans = (p_row, p_col, *chunk_plans)
                n_streams = int(os.environ.get("GBLUR_STAGED_STREAMS", world - 1))  # one per peer unless overridden
                copy_streams = [torch.cuda.Stream(device=dev) for _ in range(max(1, min(n_streams, world - 1)))]
                order = [(rank + 1 + i) % world for i in range(world - 1)]  # spread the first pushes over the links
                pad = p_col.pad

                def mirror_slots(lo_row: int, hi_row: int):
                    """reflect_101 mirror slots written by image rows [lo_row, hi_row) (problems.cuh make_blocked_dst)."""
                    out = []
                    lo, hi = max(lo_row, 1), min(hi_row - 1, pad)
                    if lo <= hi:
                        out.append((pad - hi, pad - lo + 1))
                    lo, hi = max(lo_row, rows - 1 - pad), min(hi_row - 1, rows - 2)
                    if lo <= hi:
                        out.append((pad + 2 * rows - 2 - hi, pad + 2 * rows - 2 - lo + 1))
                    return out

                def push(slot_a: int, slot_b: int, ev):
                    for i, h in enumerate(order):
                        st = copy_streams[i % len(copy_streams)]
                        st.wait_event(ev)
                        p_col.panel_push(local[h], peers[h], slot_a, slot_b, apply_to_alpha, stream=st)

                # The chunks are independent launches: issued on alternating compute streams, the ramp-down of chunk k
                # (a persistent grid draining its last line units) overlaps the ramp-up of chunk k + 1. One stream made
                # the chunked pass cost 1.22 ms against 0.85 ms for a single launch (profiles/r01/scale).
                n_comp = max(1, int(os.environ.get("GBLUR_STAGED_COMPUTE_STREAMS", "2")))
                comp_streams = [torch.cuda.Stream(device=dev) for _ in range(n_comp)] if n_comp > 1 else []

                def staged_rowpass(band: torch.Tensor):
                    comp = torch.cuda.current_stream()
                    for cs in comp_streams:
                        cs.wait_stream(comp)
                    for k, pk in enumerate(chunk_plans):
                        a, b = shard_range(r1 - r0, K, k)
                        cs = comp_streams[k % len(comp_streams)] if comp_streams else comp
                        if hybrid and (K - 1 - k) % 2 == 0:  # the last chunk goes direct: no push tail after it
                            pk.rowpass_peer(band[a:b], peers, r0 + a, apply_to_alpha, stream=cs)
                            continue
                        pk.rowpass_peer(band[a:b], local, r0 + a, apply_to_alpha, stream=cs)
                        ev = torch.cuda.Event()
                        ev.record(cs)
                        push(pad + r0 + a, pad + r0 + b, ev)
                        for m0, m1 in mirror_slots(r0 + a, r0 + b):
                            push(m0, m1, ev)
                    for cs in comp_streams:
                        comp.wait_stream(cs)
                    mark()  # the band's row pass is done; pushes of the last chunks are still in flight
                    for st in copy_streams:
                        comp.wait_stream(st)

            def fused_run(band: torch.Tensor) -> torch.Tensor:
                self.marks = []
                mark()
                fence()  # every rank has finished the previous column pass
                mark()
                if staged:
                    staged_rowpass(band)
                else:
                    p_row.rowpass_peer(band, peers, r0, apply_to_alpha)
                mark()
                fence()  # every rank's row pass (and its peer stores) has completed
                mark()
                out = torch.empty((rows, c1 - c0, channels), dtype=torch.uint8, device=dev)
                if cp < channels:
                    out[..., channels - 1] = self._exchange_alpha(band)
                p_col.colpass_panel(panel_ptr, out, (c1 - c0) * channels, apply_to_alpha)
                mark()
                return out

            self.fused_run = fused_run
            self.exchange_kind = ((f"hybrid: row pass in {K} chunks, alternately peer stores and " if hybrid else
                                   f"staged: local row pass in {K} chunks, ") +
                                  "copy-engine pushes into the owners' panel buffers (NVLink), no collective on the "
                                  "data path") if staged else \
                "fused peer stores (NVLink), no collective on the data path"
        else:
            self.exchange_kind = "NCCL all-to-all"
        return self

    def phase_ms(self):
        """After a profiled fused run (and a synchronize): ms spent in fence, row pass, fence, column pass."""
        m = self.marks
        names = ["fence_before", "row_pass_peer_stores", "fence_after", "col_pass"]
        if len(m) == 6:  # staged exchange: the row pass is local, the tail is what the copy engines still owe
            names = ["fence_before", "row_pass_local_chunks", "push_tail", "fence_after", "col_pass"]
        return {n: m[i].elapsed_time(m[i + 1]) for i, n in enumerate(names)} if len(m) == len(names) + 1 else {}

    def _exchange_alpha(self, band: torch.Tensor) -> torch.Tensor:
        """The un-blurred alpha bytes of all rows for my column band (u8 all-to-all)."""
        c0, c1 = self.my_cols
        a_send = [band[:, a:b, self.C - 1].contiguous() for a, b in self.col_ranges]
        alpha_src = torch.empty((self.rows, c1 - c0), dtype=band.dtype, device=band.device)
        exchange(a_send, [alpha_src[a:b] for a, b in self.row_ranges], self.group)
        return alpha_src

    def close(self):
        if getattr(self, "_opened", None):
            import libgaussianblur_b200 as gb

            torch.cuda.synchronize()
            if dist.is_initialized():
                dist.barrier(self.group)
            for q in self._opened:
                gb.ipc_close(q)
            self._opened = []
        for p in getattr(self, "plans", ()):
            p.close()


class HaloShardedBlur:
    """One image sharded by row bands for BOTH passes (SURVEY.md 8e third row; include/gblur_capi.h gblur_halo_*).

    Rank g holds the input rows [row_ranges[g]) — any contiguous partition — and ends up with the blurred rows
    [out_rows) of the block-aligned partition the kernels want. Per run it
      1. gathers the input rows it needs beyond its own (`pad` rows on either side, plus the few rows by which the two
         partitions differ) from the ranks that hold them: point-to-point u8 copies, a few MB;
      2. adds the column sums of its output rows into a cols*channels int32 vector and all-reduces it (the reference's
         Nyquist term along columns needs the alternating sum over the WHOLE column, src/gaussianblur.cpp:62-79);
      3. runs its row pass and its column blocks locally.
    The fp32 intermediate never leaves the GPU: about 20x fewer NVLink bytes than the all-to-all of BandShardedBlur.
    `fetch_rows(lo, hi)` / `colsum` / `run` are injectable so the plumbing is testable on CPU under gloo."""

    def __init__(self, rows: int, cols: int, channels: int, out_rows: Tuple[int, int], in_rows: Tuple[int, int],
                 colsum: Callable, run: Callable, group=None, row_ranges=None):
        self.rows, self.cols, self.C = rows, cols, channels
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.row_ranges = row_ranges or shard_ranges(rows, self.world)
        self.out_rows, self.in_rows = out_rows, in_rows
        self._colsum, self._run = colsum, run
        # what every rank needs, known to all (the partition is a pure function of the geometry)
        need = [None] * self.world
        if self.world > 1:
            dist.all_gather_object(need, tuple(in_rows), group=group)
        else:
            need = [tuple(in_rows)]
        self.need = need
        self.exchanged_bytes = 0

    def _assemble(self, band: torch.Tensor) -> torch.Tensor:
        """rows [in_rows) of the image on this rank: own rows + pieces received from the ranks that hold the rest"""
        lo, hi = self.in_rows
        W = self.cols * self.C
        my0, my1 = self.row_ranges[self.rank]
        buf = torch.empty((max(hi - lo, 0), W), dtype=band.dtype, device=band.device)
        flat = band.reshape(my1 - my0, W)
        ops, self.exchanged_bytes = [], 0
        for h in range(self.world):
            h0, h1 = self.row_ranges[h]
            # rows rank h needs from me
            n0, n1 = self.need[h]
            a, b = max(n0, my0), min(n1, my1)
            if h != self.rank and a < b:
                piece = flat[a - my0:b - my0].contiguous()
                ops.append(dist.P2POp(dist.isend, piece, h if self.group is None else dist.get_global_rank(self.group, h), self.group))
                self.exchanged_bytes += piece.numel()
            # rows I need from rank h
            a, b = max(lo, h0), min(hi, h1)
            if a < b:
                if h == self.rank:
                    buf[a - lo:b - lo].copy_(flat[a - my0:b - my0])
                else:
                    ops.append(dist.P2POp(dist.irecv, buf[a - lo:b - lo], h if self.group is None else dist.get_global_rank(self.group, h), self.group))
        if ops:
            for req in dist.batch_isend_irecv(ops):
                req.wait()
        return buf

    def run(self, band: torch.Tensor) -> torch.Tensor:
        """band: this rank's rows [row_ranges[rank]) of the image, u8. Returns the blurred rows [out_rows)."""
        W = self.cols * self.C
        rows_in = self._assemble(band)
        u = torch.zeros(W, dtype=torch.int32, device=band.device)
        self._colsum(rows_in, u)
        if self.world > 1:
            dist.all_reduce(u, group=self.group)
        out = torch.empty((max(self.out_rows[1] - self.out_rows[0], 0), self.cols, self.C), dtype=band.dtype, device=band.device)
        self._run(rows_in, u, out)
        return out

    # ---- the fast path: the input partition IS the output partition (out_ranges), buffers persist ----
    def own_view(self) -> torch.Tensor:
        """[out rows, cols, C] view into the persistent input buffer: fill it, then call run_aligned()"""
        o0, o1 = self.out_rows
        i0 = self.in_rows[0]
        return self.buf[o0 - i0:o1 - i0].view(o1 - o0, self.cols, self.C)

    def run_aligned(self) -> torch.Tensor:
        """The rank's rows [out_rows) are in own_view(). Three streams: the halo rows travel (NCCL send/recv) and the
        column sums are formed and all-reduced WHILE the row pass of the own rows runs; then the row pass of the halo
        rows, the column terms and the column blocks. Returns the persistent output buffer."""
        plan, buf = self.plan, self.buf
        (o0, o1), (i0, i1) = self.out_rows, self.in_rows
        main = torch.cuda.current_stream()
        if o1 <= o0:  # an idle rank still takes part in the collectives
            with torch.cuda.stream(self.side):
                self.side.wait_stream(main)
                self.u.zero_()
                self._halo_exchange()
                if self.world > 1:
                    dist.all_reduce(self.u, group=self.group)
            main.wait_stream(self.side)
            return self.out
        prof = getattr(self, "profile", False)
        marks = {}

        def mark(name, stream):
            if prof:
                e = torch.cuda.Event(enable_timing=True)
                e.record(stream)
                marks[name] = e

        mark("start", main)
        with torch.cuda.stream(self.xs):
            self.xs.wait_stream(main)
            self._halo_exchange()
            mark("halo_rows_arrived", self.xs)
        with torch.cuda.stream(self.side):
            self.side.wait_stream(main)
            self.u.zero_()
            plan.colsum(buf, self.u, stream=self.side)
            mark("colsum_done", self.side)
        plan.rowpass(buf, o0, o1)
        mark("rowpass_own_done", main)
        # the all-reduce is enqueued after the send/recv on every rank (one communicator: same order everywhere)
        with torch.cuda.stream(self.side):
            self.side.wait_stream(self.xs)
            if self.world > 1:
                dist.all_reduce(self.u, group=self.group)
            mark("allreduce_done", self.side)
        main.wait_stream(self.xs)
        plan.rowpass2(buf, i0, o0, o1, i1)  # the halo rows above and below in one launch
        mark("rowpass_halo_done", main)
        main.wait_stream(self.side)
        mark("terms_ready", main)
        plan.colpass(self.u, self.out)
        mark("colpass_done", main)
        if prof:
            self._marks = marks
        return self.out

    def phase_ms(self) -> dict:
        """after a run_aligned() with self.profile = True: when each point of the three streams was reached, in ms from the
        start of the call (device clocks)"""
        torch.cuda.synchronize()
        m = self._marks
        return {"t_" + k: m["start"].elapsed_time(v) for k, v in m.items() if k != "start"}

    def _halo_exchange(self):
        """pad rows above and below from the ranks that own them (out_ranges partition), into the persistent buffer"""
        (i0, i1) = self.in_rows
        my0, my1 = self.out_rows
        ops, self.exchanged_bytes = [], 0
        for h in range(self.world):
            if h == self.rank:
                continue
            h0, h1 = self.out_ranges[h]
            n0, n1 = self.need[h]
            a, b = max(n0, my0), min(n1, my1)  # rows rank h needs from me
            peer = h if self.group is None else dist.get_global_rank(self.group, h)
            if a < b and n1 > n0:
                ops.append(dist.P2POp(dist.isend, self.buf[a - i0:b - i0], peer, self.group))
                self.exchanged_bytes += (b - a) * self.cols * self.C
            a, b = max(i0, h0), min(i1, h1)    # rows I need from rank h
            if a < b and my1 > my0:
                ops.append(dist.P2POp(dist.irecv, self.buf[a - i0:b - i0], peer, self.group))
        if ops:
            for req in dist.batch_isend_irecv(ops):
                req.wait()

    @classmethod
    def cuda(cls, rows: int, cols: int, channels: int, sigma: float, device: int = 0, group=None) -> "HaloShardedBlur":
        import libgaussianblur_b200 as gb

        world = dist.get_world_size(group) if dist.is_initialized() else 1
        rank = dist.get_rank(group) if dist.is_initialized() else 0
        plan = gb.HaloPlan(rows, cols, channels, sigma, world, rank, device=device)
        self = cls(rows, cols, channels, (plan.out_row0, plan.out_row1), (plan.in_row0, plan.in_row1),
                   colsum=lambda rows_in, u: plan.colsum(rows_in, u) if plan.out_row1 > plan.out_row0 else None,
                   run=lambda rows_in, u, out: plan.run(rows_in, u, out) if plan.out_row1 > plan.out_row0 else None,
                   group=group)
        self.plan = plan
        dev = torch.device("cuda", device)
        parts = [gb.halo_partition(rows, cols, sigma, world, r) for r in range(world)]
        self.out_ranges = [(q["out_row0"], q["out_row1"]) for q in parts]
        W = cols * channels
        self.buf = torch.zeros((max(plan.in_row1 - plan.in_row0, 1), W), dtype=torch.uint8, device=dev)
        self.out = torch.empty((max(plan.out_row1 - plan.out_row0, 0), cols, channels), dtype=torch.uint8, device=dev)
        self.u = torch.zeros(W, dtype=torch.int32, device=dev)
        self.side, self.xs = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
        self.exchange_kind = ("halo: row bands for both passes, pad input rows from the neighbours + one int32 all-reduce "
                              "(NCCL), no exchange of the fp32 intermediate")
        return self

    def close(self):
        if getattr(self, "plan", None) is not None:
            self.plan.close()
            self.plan = None


class FrameShardedBlur:
    """A batch of equal-geometry frames sharded by contiguous frame blocks; no collective. Every rank
    builds the same plan once (kernel spectrum and twiddles reused for all its frames)."""

    def __init__(self, n_frames: int, rows: int, cols: int, channels: int, sigma: float, device: int = 0,
                 chunk_frames: int = 16, group=None):
        import libgaussianblur_b200 as gb

        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.frames = shard_range(n_frames, self.world, self.rank)
        self.chunk = max(1, min(chunk_frames, self.frames[1] - self.frames[0]))
        self.plan = gb.Plan(rows, cols, channels, sigma, device=device, max_frames=self.chunk)

    def run(self, d_frames: torch.Tensor, d_out: torch.Tensor | None = None, apply_to_alpha: bool = False):
        """d_frames: this rank's frames [n_local, rows, cols, C] u8 on the device."""
        d_out = d_frames if d_out is None else d_out
        n = d_frames.shape[0]
        for f in range(0, n, self.chunk):
            k = min(self.chunk, n - f)
            self.plan.run_device(d_frames[f:f + k], d_out[f:f + k], apply_to_alpha, n_frames=k)
        return d_out

    def close(self):
        self.plan.close()
