"""TeraSort on the engine: the flagship workload (BASELINE.json config 2).

Phases::

    setup   teragen on device -> MapTask per map (sort + partition + IFile/CRC) -> MOFs in HBM
            -> all-gather index tables -> symmetric arena -> push descriptors / run tables
    step    [timed]  fused shuffle push (every MOF partition -> its reducer's arena over NVLink)
                     -> landed-flag acquire -> key normalise -> k-way merge -> record gather
    check   TeraValidate equivalent: order inside every reducer, key range per reducer, record
            count and order-independent fingerprint equal to the generated input.

One reducer per rank, one process per GPU (the reference has the same limit:
"we only support 1 reducer per process", src/Merger/reducer.h:137).
"""
from __future__ import annotations

import os
from typing import List, Optional

import numpy as np
import torch
import torch.distributed as dist

from ..mof import TERA_STRIDE, DeviceMOF, MapTask, uniform_splitters_w0
from ..ops import kernels as K
from ..ops.merger import DeviceMerger
from ..parallel.arena import SymmetricArena
from ..parallel.shuffle import ShufflePlan


def _host_mem_available():
    """MemAvailable from /proc/meminfo, capped by the cgroup limit when there is one (bytes)."""
    avail = None
    try:
        with open("/proc/meminfo") as f:
            for line in f:
                if line.startswith("MemAvailable:"):
                    avail = int(line.split()[1]) * 1024
                    break
        for path in ("/sys/fs/cgroup/memory.max", "/sys/fs/cgroup/memory/memory.limit_in_bytes"):
            try:
                v = open(path).read().strip()
                if v.isdigit() and int(v) < (1 << 60):
                    avail = min(avail, int(v)) if avail is not None else int(v)
            except OSError:
                pass
    except OSError:
        return None
    return avail


class TeraSortJob:
    #: bytes of one slot of the pinned host ring that receives the merged stream
    E2E_RING_BYTES = 256 << 20

    def __init__(self, records_per_rank: int, maps_per_rank: int, device: Optional[torch.device] = None, seed: int = 1234,
                 key_bits: int = 0, group=None, push_ctas: int = 0, subparts: int = 1, merge_ctas_per_sm: int = 0,
                 push_mode: str = "lsu", local_zero_copy: bool = False, tma_flush: int = 0, local_copy: str = "auto", arena_slots: int = 0,
                 device_cuts: bool = False):
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.group = group
        self.n_local = records_per_rank
        self.maps = maps_per_rank
        self.seed, self.key_bits = seed, key_bits
        self.push_ctas = push_ctas
        #: every reducer's key range is cut into `subparts` consecutive sub-partitions that are pushed in order on a
        #: side stream while the merge of the previous sub-partition runs: the NVLink-bound (or, on one GPU,
        #: HBM-bound) push overlaps the issue-bound merge. 1 = push everything, then merge.
        self.subparts = max(1, subparts)
        self.merge_ctas_per_sm = merge_ctas_per_sm
        #: "lsu": load/store push kernel on every SM (one launch per sub-partition); "tma": copy-engine push, ONE
        #: launch for all sub-partitions on `push_ctas` SMs of its own, the merge keeps the remaining SMs
        assert push_mode in ("lsu", "tma")
        self.push_mode = push_mode
        #: copy-engine push: chunks between two advances of the landed-byte counters (0 = library default)
        self.tma_flush = tma_flush
        #: merge this rank's own map outputs where they lie instead of copying them into the arena first (the
        #: reference and Hadoop copy local fetches like remote ones). Off by default: the headline benchmark moves
        #: every segment. Not available on the end-to-end path, whose map outputs alternate between two buffers.
        self.local_zero_copy = local_zero_copy
        #: how a rank's OWN map outputs reach its arena. "push": through the push kernel like the remote ones. "ce": with
        #: the GPU's copy engines (one device-to-device cudaMemcpyAsync per segment on a stream of its own, then one
        #: atomic add of the landed bytes per sub-partition), so the SMs reserved for the push move only what must cross
        #: NVLink. "auto": "ce" when the step is pipelined (several sub-partitions), "push" otherwise.
        assert local_copy in ("auto", "push", "ce")
        self.local_copy = local_copy
        #: receive slots of the arena (0 = one per sub-partition: the arena holds the whole partition). Fewer slots than
        #: sub-partitions make the arena a ring with credit-based back-pressure: receive memory = slots / subparts of the
        #: partition (copy-engine push only).
        self.arena_slots = arena_slots
        self.ring_steps = 0
        #: False: the map side writes P x subparts partitions (uniform sub-ranges), one segment per sub-partition.
        #: True: the map outputs have the P partitions Hadoop writes (what src/MOFServer/IndexInfo.cc:237-335 serves) and the
        #: sub-partitions are cut on the device: splitters sampled from the sorted segments of all ranks (equi-depth, so skewed
        #: keys still give balanced sub-partitions), then one lower-bound search per (segment, splitter).
        self.device_cuts = device_cuts
        #: polls of the landed-byte counters before the wait kernel gives up and records a time-out in the arena's
        #: status word (the watchdog that turns a lost push into a failure up-call instead of a hang)
        self.wait_max_spins = 50_000_000
        self.mofs: List[DeviceMOF] = []
        self.input_fingerprint = 0
        self.steps_run = 0

    # ------------------------------------------------------------------------------------ setup
    def setup(self) -> "TeraSortJob":
        dev, P = self.device, self.world
        per_map = [self.n_local // self.maps + (1 if i < self.n_local % self.maps else 0) for i in range(self.maps)]
        C = self.subparts
        cmap = 1 if self.device_cuts else C
        splitters = uniform_splitters_w0(P * cmap)
        task = MapTask(max(per_map), P * cmap, dev)
        row0 = self.rank * self.n_local
        fp = 0
        for m, nm in enumerate(per_map):
            recs = K.teragen(nm, self.seed, first_row=row0, framed=True, key_bits=self.key_bits, device=dev)
            _bad, f, cnt = K.validate_fixed(recs, nm, TERA_STRIDE, 2)
            assert cnt == nm
            fp = (fp + f) & 0xFFFFFFFFFFFFFFFF
            self.mofs.append(task.run(recs, nm, splitters))
            row0 += nm
            del recs
        self.input_fingerprint = fp
        del task
        torch.cuda.synchronize(dev)

        # ---- index exchange (plumbing over the process group)
        local_idx = self._device_cut_index() if (self.device_cuts and C > 1) else np.stack([m.index for m in self.mofs])  # [maps, P*C, 3]
        if P > 1:
            t = torch.from_numpy(local_idx).to(dev)
            allt = torch.empty((P,) + tuple(t.shape), dtype=torch.int64, device=dev)
            dist.all_gather_into_tensor(allt, t, group=self.group)
            index = allt.cpu().numpy()
        else:
            index = local_idx[None]
        self.plan = ShufflePlan(index, record_stride=TERA_STRIDE, subparts=C, slots=self.arena_slots)
        if self.plan.slots < C:
            assert self.push_mode == "tma" and not self.local_zero_copy, "the receive ring needs the copy-engine push"
        nruns = self.plan.nruns

        # ---- arena + tables (one push table and one run table per sub-partition)
        self.arena = SymmetricArena(self.plan.max_arena_bytes(), nruns * C, dev, self.group)
        if self.local_copy == "auto":
            self.local_copy = "ce" if (C > 1 and not self.local_zero_copy) else "push"
        if self.local_zero_copy:
            self.local_copy = "push"
        self.local_stream = torch.cuda.Stream(dev) if self.local_copy == "ce" else None
        self.ev_local, self.ev_merged = torch.cuda.Event(), torch.cuda.Event()
        self.push_set = self._push_tables([m.data for m in self.mofs])
        self.push_descs_dev = self.push_set["tables"]
        self.nseg = self.maps * P
        self.push_scratch = [torch.empty(K.push_scratch_bytes(self.nseg), dtype=torch.uint8, device=dev) for _ in range(C)]
        self.push_descs_all = self.push_set["all"]
        self.push_tma_scratch = torch.empty(K.push_tma_scratch_bytes(self.nseg * C), dtype=torch.uint8, device=dev)
        self.sub_runs, self.sub_n, expect = [], [], []
        max_sub = 1
        for j in range(C):
            rd, run_off, run_base, run_bytes = self.plan.run_descs(self.rank, self.arena.peer_data_ptr(self.rank), j)
            if self.local_zero_copy:
                for m, mof in enumerate(self.mofs):
                    rid = self.plan.run_id(self.rank, m)
                    base = mof.data.data_ptr() + int(self.plan.index[self.rank, m, self.rank * C + j, 0])
                    rd["base"][rid] = base
                    run_base[rid] = base
                    run_bytes[rid] = 0                       # nothing will land for this run
            self.sub_n.append(int(run_off[-1]))
            max_sub = max(max_sub, self.sub_n[-1])
            expect.append(run_bytes.astype(np.int64))
            self.sub_runs.append((rd, run_off, run_base))
        self.n_out = sum(self.sub_n)
        self.expect_step = torch.from_numpy(np.concatenate(expect)).to(dev)
        self.expect = torch.zeros_like(self.expect_step)
        self.merger = DeviceMerger(max_sub, nruns, TERA_STRIDE, dev, out_bytes=self.n_out * TERA_STRIDE + 2,
                                   ctas_per_sm=self.merge_ctas_per_sm)
        snaps = []
        for rd, run_off, run_base in self.sub_runs:
            self.merger.set_runs(rd, run_off, run_base)
            snaps.append(self.merger.snapshot_runs())
        self.sub_runs = snaps
        self.push_stream = torch.cuda.Stream(dev) if C > 1 else None
        self.ev_go, self.ev_pushed = torch.cuda.Event(), torch.cuda.Event()
        self.bytes_in = int(sum(int(m.index[:, 2].sum()) for m in self.mofs))
        self.record_bytes_local = self.n_local * 100
        self._warm_up()
        torch.cuda.synchronize(dev)
        if P > 1:
            dist.barrier(group=self.group)
        return self

    def _device_cut_index(self, samples_per_segment: int = 64) -> np.ndarray:
        """Index ``[maps, P * subparts, 3]`` of sub-segments cut on the device out of P-partition map outputs.

        Every rank samples ``samples_per_segment`` equi-depth keys from each of its sorted segments; the samples (and segment
        sizes) are all-gathered, and every rank derives the same ``subparts - 1`` splitters per reducer as weighted quantiles of
        that reducer's samples. The cut positions are lower bounds of the splitters inside each segment
        (uda_lower_bound_records). A sub-segment is a slice of its parent: body bytes only, no EOF marker or checksum of its own."""
        dev, P, C, q = self.device, self.world, self.subparts, samples_per_segment
        maps = len(self.mofs)
        samples = torch.zeros((maps, P, q), dtype=torch.int64, device=dev)
        nrec = np.zeros((maps, P), dtype=np.int64)
        for m, mof in enumerate(self.mofs):
            for p in range(P):
                start, raw, _part = (int(x) for x in mof.index[p])
                nrec[m, p] = (raw - 2) // TERA_STRIDE
                if nrec[m, p] > 0:
                    samples[m, p] = K.sample_splitters(mof.data[start:], int(nrec[m, p]), TERA_STRIDE, 2, 10, q)
        nrec_dev = torch.from_numpy(nrec).to(dev)
        if P > 1:
            all_s = torch.empty((P,) + tuple(samples.shape), dtype=torch.int64, device=dev)
            all_n = torch.empty((P,) + tuple(nrec_dev.shape), dtype=torch.int64, device=dev)
            dist.all_gather_into_tensor(all_s, samples.contiguous(), group=self.group)
            dist.all_gather_into_tensor(all_n, nrec_dev, group=self.group)
        else:
            all_s, all_n = samples[None], nrec_dev[None]
        hs = all_s.cpu().numpy().view(np.uint64)          # [world, maps, P, q]
        hn = all_n.cpu().numpy().astype(np.float64)       # [world, maps, P]
        split = np.zeros((P, C - 1), dtype=np.uint64)
        for p in range(P):
            keys = hs[:, :, p, :].reshape(-1)
            w = np.repeat(hn[:, :, p].reshape(-1) / q, q)
            order = np.argsort(keys, kind="stable")
            cum = np.cumsum(w[order])
            total = cum[-1] if len(cum) else 0.0
            for j in range(1, C):
                i = int(np.searchsorted(cum, total * j / C, side="left")) if total > 0 else 0
                split[p, j - 1] = keys[order][min(i, len(keys) - 1)] if len(keys) else 0
            split[p] = np.maximum.accumulate(split[p])
        self.sub_splitters = split
        split_dev = torch.from_numpy(split.view(np.int64)).to(dev)
        index = np.zeros((maps, P * C, 3), dtype=np.int64)
        for m, mof in enumerate(self.mofs):
            for p in range(P):
                start, _raw, _part = (int(x) for x in mof.index[p])
                n = int(nrec[m, p])
                cuts = K.lower_bound_records(mof.data[start:], n, TERA_STRIDE, 2, 10, split_dev[p]).cpu().numpy() if n > 0 else np.zeros(C - 1, np.int64)
                b = np.concatenate([[0], np.minimum(np.maximum.accumulate(cuts), n), [n]]).astype(np.int64)
                for j in range(C):
                    body = int(b[j + 1] - b[j]) * TERA_STRIDE
                    index[m, p * C + j] = (start + int(b[j]) * TERA_STRIDE, body + 2, body)
        return index

    def _warm_up(self) -> None:
        """Runs every kernel a step launches once, on an idle device. Kernels are loaded lazily at their first launch, and
        that load waits for the kernels already running: inside a step some of those spin on the progress of kernels that are
        launched later (the flag wait on the push, the push on the ring credits, which come from the merges), so a first
        launch in the middle of a step blocks until a watchdog gives up. (Measured: the first ring step took 590 ms and
        ended with a push time-out; the following steps 1.5 ms.) Nothing here touches the arena, counters or credits."""
        dev, P, C = self.device, self.world, self.subparts
        k = self.plan.nruns
        # a tiny merge with the job's run count (selects the same partition kernel) over throw-away records
        nrec = 4
        recs = K.teragen(nrec * k, self.seed + 99, first_row=0, framed=True, key_bits=self.key_bits, device=dev)
        pad = torch.zeros(recs.numel() + 256, dtype=torch.uint8, device=dev)
        pad[64:64 + recs.numel()] = recs          # unsorted runs are fine: only the launches matter
        desc, _off = K.make_run_descs([pad.data_ptr() + 64 + r * nrec * TERA_STRIDE for r in range(k)], [nrec] * k)
        tmp_out = torch.empty(nrec * k * TERA_STRIDE + 66, dtype=torch.uint8, device=dev)
        scratch = torch.empty(K.merge_records_scratch_bytes(nrec * k, k), dtype=torch.uint8, device=dev)
        if self.merger.fused:
            K.merge_records_fixed(K.to_device_struct(desc, dev), k, nrec * k, nrec, TERA_STRIDE, 2, 10, tmp_out, scratch, write_eof=True,
                                  align2=True, ctas_per_sm=self.merge_ctas_per_sm)
        # push kernels over a table of empty descriptors
        empty = self.push_descs_all.clone()
        host = empty.cpu().numpy().view(K.PUSH_DESC_DTYPE).copy()
        host["bytes"][:] = 0
        empty.copy_(torch.from_numpy(host.view(np.uint8).reshape(-1)))
        dummy = torch.zeros(128, dtype=torch.int64, device=dev)
        status = torch.zeros(4, dtype=torch.int32, device=dev)
        if self.push_mode == "tma":
            K.shuffle_push_tma(empty, self.nseg * C, C, P, self.push_tma_scratch, max(1, self.push_ctas), self.tma_flush)
            if self.plan.slots < C:
                K.shuffle_push_tma_ring(empty, self.nseg * C, C, P, self.push_tma_scratch, max(1, self.push_ctas), self.tma_flush, dummy, P, -C,
                                        status, 1000)
        K.shuffle_push(empty[: self.push_descs_dev[0].numel()], self.nseg, self.push_scratch[0], 8, P)
        # flag wait (already satisfied), credit publication into a dummy table, the torch kernels of the copy-engine path
        K.wait_flags(dummy[:4], dummy[4:8], 4, status, 10)
        slots = torch.tensor([dummy.data_ptr() + 8 * (16 + p) for p in range(P)], dtype=torch.int64, device=dev)
        K.credit_publish(slots, P, dummy[8:9])
        dummy.index_add_(0, torch.tensor([40, 41], dtype=torch.int64, device=dev), torch.tensor([0, 0], dtype=torch.int64, device=dev))
        dummy[64:96].add_(dummy[96:128])
        tmp_out[:64].copy_(pad[:64], non_blocking=True)
        if P > 1:
            self.arena.barrier()      # collective: every rank calls setup()
        torch.cuda.synchronize(dev)

    # ------------------------------------------------------------------------------------ step
    def _push_tables(self, mof_data: List[torch.Tensor]) -> dict:
        """Everything one set of map-output buffers needs for a step: the push descriptor table of every sub-partition
        (``tables``, and ``all`` = their concatenation for the one-launch copy-engine push) and, with ``local_copy ==
        "ce"``, the device-to-device copies of this rank's own segments (``local[j]`` = list of (dst view, src view),
        ``local_idx[j]`` / ``local_bytes[j]`` = counter slots and landed bytes)."""
        P, C = self.world, self.subparts
        mof_ptrs = [t.data_ptr() for t in mof_data]
        tables, local, local_idx, local_bytes = [], [], [], []
        arena = self.arena.data()
        for j in range(C):
            d = self.plan.push_descs(self.rank, mof_ptrs, [self.arena.peer_data_ptr(p) for p in range(P)],
                                     [self.arena.peer_flag_ptr(p, 0) for p in range(P)], j)
            if self.local_zero_copy or self.local_copy == "ce":
                d["bytes"][(P - 1) * self.maps:] = 0      # the last `maps` descriptors are the ones addressed to this rank
            tables.append(K.to_device_struct(d, self.device))
            if self.local_copy == "ce":
                lay = self.plan.layout(self.rank, j)
                jobs, idx, nb = [], [], []
                for m in range(self.maps):
                    start, _raw, part = (int(x) for x in self.plan.index[self.rank, m, self.rank * C + j])
                    rid = self.plan.run_id(self.rank, m)
                    if part > 0:
                        off = int(lay.run_offsets[rid])
                        jobs.append((arena[off:off + part], mof_data[m][start:start + part]))
                    idx.append(j * self.plan.nruns + rid)
                    nb.append(part)
                local.append(jobs)
                local_idx.append(torch.tensor(idx, dtype=torch.int64, device=self.device))
                local_bytes.append(torch.tensor(nb, dtype=torch.int64, device=self.device))
        return {"tables": tables, "all": torch.cat(tables).contiguous(), "local": local, "local_idx": local_idx, "local_bytes": local_bytes}

    def _local_copies(self, pset: dict, subs) -> None:
        """Queues this rank's own segments of sub-partitions ``subs`` on the copy engines (stream of their own); the
        landed-byte counters of a sub-partition advance after its copies (stream order), with atomics on those slots only."""
        flags = self.arena.flags()
        with torch.cuda.stream(self.local_stream):
            for j in subs:
                for dst, src in pset["local"][j]:
                    dst.copy_(src, non_blocking=True)
                flags.index_add_(0, pset["local_idx"][j], pset["local_bytes"][j])
            self.ev_local.record(self.local_stream)

    def _merge_sub(self, j: int, out: torch.Tensor, tl: Optional[dict] = None) -> None:
        C, nruns = self.subparts, self.plan.nruns
        K.wait_flags(self.arena.flags()[j * nruns:(j + 1) * nruns], self.expect[j * nruns:(j + 1) * nruns], nruns, self.arena.status,
                     self.wait_max_spins)
        if tl is not None:
            tl["m0"][j].record(torch.cuda.current_stream(self.device))
        self.merger.use_runs(self.sub_runs[j])
        off = sum(self.sub_n[:j]) * TERA_STRIDE
        self.merger.merge_fixed(out=out[off:], write_eof=(j == C - 1))
        if tl is not None:
            tl["m1"][j].record(torch.cuda.current_stream(self.device))

    def step(self, pset: Optional[dict] = None, out: Optional[torch.Tensor] = None, _timeline: Optional[dict] = None) -> torch.Tensor:
        """One shuffle + merge. Flags are cumulative counters, so step ``i`` waits for ``i * bytes``; the
        all-rank barrier at the top keeps a fast rank from overwriting an arena that is still being merged.
        With ``subparts > 1`` the pushes of all sub-partitions are queued on a side stream and the merge of
        sub-partition j starts as soon as its runs have landed, i.e. while sub-partition j+1 is still in flight.
        With ``arena_slots < subparts`` the arena is a ring: sub-partition j lands in slot ``j % slots`` after the
        merge of sub-partition ``j - slots`` (credits for the remote senders, stream order for the local copies).
        ``pset``: the tables of another set of map-output buffers (:meth:`_push_tables`), default the job's own."""
        self.steps_run += 1
        pset = self.push_set if pset is None else pset
        descs = pset["tables"]
        out = self.merger.out if out is None else out
        C, S = self.subparts, self.plan.slots
        ring = S < C
        cur = torch.cuda.current_stream(self.device)
        if self.world > 1:
            self.arena.barrier()
        self.expect.add_(self.expect_step)
        tma = self.push_mode == "tma"
        ce = self.local_copy == "ce"
        # one GPU, copy engines: the push kernel would find only empty descriptors (UDA_FORCE_PUSH_KERNEL=1 launches it anyway: tests)
        nothing_remote = ce and self.world == 1 and not os.environ.get("UDA_FORCE_PUSH_KERNEL")
        tl = _timeline
        if tl is not None:
            tl["t0"].record(cur)
        self.ev_go.record(cur)
        if ce:
            self.local_stream.wait_event(self.ev_go)
            self._local_copies(pset, range(min(S, C)))
        push_stream = self.push_stream if C > 1 else cur
        with torch.cuda.stream(push_stream):
            if C > 1:
                push_stream.wait_event(self.ev_go)
            if nothing_remote:
                pass
            elif tma and ring:
                # cumulative sub-partitions merged before this step = ring_steps * C; sub-partition j needs j + 1 - S more
                K.shuffle_push_tma_ring(pset["all"], self.nseg * C, C, self.world, self.push_tma_scratch, self.push_ctas, self.tma_flush,
                                        self.arena.credits, self.world, self.ring_steps * C + 1 - S, self.arena.status, self.wait_max_spins)
            elif tma:
                K.shuffle_push_tma(pset["all"], self.nseg * C, C, self.world, self.push_tma_scratch, self.push_ctas, self.tma_flush)
            else:
                for j in range(C):
                    K.shuffle_push(descs[j], self.nseg, self.push_scratch[j], self.push_ctas, self.world)
                    if tl is not None:
                        tl["pushed"][j].record(push_stream)
            if tl is not None and (tma or nothing_remote):
                for j in range(C):
                    tl["pushed"][j].record(push_stream)
            if C > 1:
                self.ev_pushed.record(push_stream)
        for j in range(C):
            self._merge_sub(j, out, tl)
            if ring:
                self.arena.publish_credit()            # slot j % S is free: remote senders may refill it ...
                if ce and j + S < C:                   # ... and so may the local copies of sub-partition j + S
                    self.ev_merged.record(cur)
                    self.local_stream.wait_event(self.ev_merged)
                    self._local_copies(pset, [j + S])
        if ring:
            self.ring_steps += 1
        if C > 1:
            cur.wait_event(self.ev_pushed)  # the step is over only when this rank's own pushes have retired
        if ce:
            cur.wait_event(self.ev_local)
        return out[: self.n_out * TERA_STRIDE + 2]

    def capture(self):
        """Captures one :meth:`step` into a CUDA graph and returns ``(graph, out)``; ``graph.replay()`` then runs a
        whole shuffle + merge with one launch from the host (16 kernels on one GPU, ~170 on eight). Nothing in a step
        depends on host state: the landed-byte expectations and the barrier epoch advance on the device. With
        sub-partition pipelining the graph has two branches (push, wait + merge) that must run concurrently, exactly
        as the two streams do in eager mode."""
        torch.cuda.synchronize(self.device)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            out = self.step()
        return g, out

    def step_timeline(self) -> dict:
        """One pipelined step with an event after every push (side stream) and around every sub-partition merge
        (main stream); times in ms since the start of the step (diagnostics for the overlap)."""
        C = self.subparts
        if C == 1:
            return {}
        mk = lambda: torch.cuda.Event(enable_timing=True)
        tl = {"t0": mk(), "pushed": [mk() for _ in range(C)], "m0": [mk() for _ in range(C)], "m1": [mk() for _ in range(C)]}
        self.step(_timeline=tl)
        torch.cuda.synchronize(self.device)
        t0 = tl["t0"]
        return {"pushed_ms": [t0.elapsed_time(e) for e in tl["pushed"]], "merge_start_ms": [t0.elapsed_time(e) for e in tl["m0"]],
                "merge_end_ms": [t0.elapsed_time(e) for e in tl["m1"]]}

    def step_phases(self) -> dict:
        """One step with CUDA events between the phases, run one after the other: everything pushed, then everything
        waited for, then everything merged (diagnostics; not used for the headline number). With a receive ring the phases
        cannot be separated (a push waits for merges): only the whole step is timed."""
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        if self.plan.slots < self.subparts:
            ev[0].record()
            self.step()
            ev[3].record()
            torch.cuda.synchronize(self.device)
            return {"step_ms": ev[0].elapsed_time(ev[3])}
        cur = torch.cuda.current_stream(self.device)
        if self.world > 1:
            self.arena.barrier()
        self.expect.add_(self.expect_step)
        ev[0].record()
        if self.local_copy == "ce":   # on the copy engines, concurrently with the push kernel below
            self.ev_go.record(cur)
            self.local_stream.wait_event(self.ev_go)
            self._local_copies(self.push_set, range(self.subparts))
        if self.local_copy == "ce" and self.world == 1:
            pass
        elif self.push_mode == "tma":
            K.shuffle_push_tma(self.push_descs_all, self.nseg * self.subparts, self.subparts, self.world, self.push_tma_scratch,
                               self.push_ctas, self.tma_flush)
        else:
            for j in range(self.subparts):  # phases are measured un-overlapped, one sub-partition after the other
                K.shuffle_push(self.push_descs_dev[j], self.nseg, self.push_scratch[j], self.push_ctas, self.world)
        if self.local_copy == "ce":
            cur.wait_event(self.ev_local)
        ev[1].record()
        nr = self.plan.nruns * self.subparts
        K.wait_flags(self.arena.flags(), self.expect, nr, self.arena.status)
        ev[2].record()
        for j in range(self.subparts):
            self._merge_sub(j, self.merger.out)
        ev[3].record()
        torch.cuda.synchronize(self.device)
        self.steps_run += 1
        return {"push_ms": ev[0].elapsed_time(ev[1]), "wait_ms": ev[1].elapsed_time(ev[2]), "merge_ms": ev[2].elapsed_time(ev[3])}

    # ------------------------------------------------------------------------------------ end to end
    def e2e_setup(self) -> None:
        """End-to-end path: map outputs start in pinned host memory (what a provider that just read
        ``file.out`` from disk holds) and the merged stream must end in pinned host memory (the buffer a
        JVM would wrap as a direct ByteBuffer, reference: src/UdaBridge.cc:535-551). Device input and
        output buffers are double-buffered so H2D of step i+1, compute of step i and D2H of step i-1
        overlap on three streams."""
        dev = self.device
        assert not self.local_zero_copy, "the end-to-end path double-buffers the map outputs: local_zero_copy is not supported there"
        out_bytes = self.n_out * TERA_STRIDE + 2
        need = sum(int(m.index[-1, 0] + m.index[-1, 2]) for m in self.mofs) + 2 * self.E2E_RING_BYTES
        avail = _host_mem_available()
        if avail is not None and need * self.world > 0.7 * avail:
            raise RuntimeError(f"not enough host memory to pin the e2e staging buffers: need {need * self.world >> 30} GiB "
                               f"across {self.world} ranks, {avail >> 30} GiB available")
        self.h_mofs = []
        for m in self.mofs:
            nb = int(m.index[-1, 0] + m.index[-1, 2])
            h = torch.empty(nb, dtype=torch.uint8, pin_memory=True)
            h.copy_(m.data[:nb])
            self.h_mofs.append(h)
        self.e2e_h2d_bytes = sum(h.numel() for h in self.h_mofs)
        self.e2e_d2h_bytes = out_bytes
        # The merged stream is consumed in blocks and never materialised on the host (the reference hands
        # Java 1 MiB blocks out of two staging buffers, src/include/NetlevComm.h:33): a two-slot pinned ring.
        self.h_ring = [torch.empty(self.E2E_RING_BYTES, dtype=torch.uint8, pin_memory=True) for _ in range(2)]
        # second device buffer set
        self.d_mofs = [[m.data for m in self.mofs], [torch.empty_like(m.data) for m in self.mofs]]
        self.d_descs = [self.push_set, self._push_tables(self.d_mofs[1])]
        self.d_out = [self.merger.out, torch.empty_like(self.merger.out)]
        self.s_h2d, self.s_d2h = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
        self.ev_h2d = [torch.cuda.Event(), torch.cuda.Event()]
        self.ev_comp = [torch.cuda.Event(), torch.cuda.Event()]
        self.ev_d2h = [torch.cuda.Event(), torch.cuda.Event()]
        self.e2e_i = 0
        torch.cuda.synchronize(dev)

    def e2e_step(self) -> torch.Tensor:
        i, b = self.e2e_i, self.e2e_i % 2
        self.e2e_i += 1
        cur = torch.cuda.current_stream(self.device)
        with torch.cuda.stream(self.s_h2d):
            if i >= 2:
                self.s_h2d.wait_event(self.ev_comp[b])      # step i-2 has finished reading this MOF set
            for h, d in zip(self.h_mofs, self.d_mofs[b]):
                d[: h.numel()].copy_(h, non_blocking=True)
            self.ev_h2d[b].record(self.s_h2d)
        cur.wait_event(self.ev_h2d[b])
        if i >= 2:
            cur.wait_event(self.ev_d2h[b])                   # step i-2's result has left this output buffer
        out = self.step(self.d_descs[b], self.d_out[b])
        self.ev_comp[b].record(cur)
        with torch.cuda.stream(self.s_d2h):
            self.s_d2h.wait_event(self.ev_comp[b])
            nb, ring = out.numel(), self.E2E_RING_BYTES
            for k, off in enumerate(range(0, nb, ring)):
                ln = min(ring, nb - off)
                self.h_ring[k % 2][:ln].copy_(out[off:off + ln], non_blocking=True)
            self.ev_d2h[b].record(self.s_d2h)
        self._e2e_last = b
        return out

    def e2e_check(self) -> bool:
        """Validate the device stream the host blocks were copied from, and that the last block that
        reached host memory is byte-identical to its device source."""
        torch.cuda.synchronize(self.device)
        out = self.d_out[self._e2e_last]
        nb, ring = self.n_out * TERA_STRIDE + 2, self.E2E_RING_BYTES
        nblk = (nb + ring - 1) // ring
        off = (nblk - 1) * ring
        last_ok = bool(torch.equal(self.h_ring[(nblk - 1) % 2][: nb - off], out[off:nb].cpu()))
        return bool(self.check(out)["ok"]) and last_ok

    #: kernel launches of this library inside one step (barrier, then per sub-partition: plan, push, wait, merge chain)
    @property
    def launches_per_step(self) -> int:
        push = 2 if self.push_mode == "tma" else 2 * self.subparts     # plan + push: one pair for all sub-partitions with tma
        if self.local_copy == "ce" and self.world == 1:
            push = 0                                                   # everything travels on the copy engines
        return (1 if self.world > 1 else 0) + push + self.subparts * (1 + DeviceMerger.LAUNCHES_PER_MERGE)

    # ------------------------------------------------------------------------------------ check
    def check(self, out: Optional[torch.Tensor] = None) -> dict:
        out = self.merger.out if out is None else out
        dev = self.device
        status = int(self.arena.status[0].item())
        bad, fp, cnt = K.validate_fixed(out, self.n_out, TERA_STRIDE, 2)
        eof_ok = bool((out[self.n_out * TERA_STRIDE: self.n_out * TERA_STRIDE + 2].cpu().numpy() == 0xFF).all())
        # key range of this reducer
        lo = 0 if self.rank == 0 else ((1 << 64) * self.rank) // self.world
        hi = (1 << 64) if self.rank == self.world - 1 else ((1 << 64) * (self.rank + 1)) // self.world
        range_ok = True
        if self.n_out:
            first = int.from_bytes(bytes(out[2:10].cpu().numpy()), "big")
            last_off = (self.n_out - 1) * TERA_STRIDE + 2
            last = int.from_bytes(bytes(out[last_off:last_off + 8].cpu().numpy()), "big")
            range_ok = lo <= first and last < hi
        tot = torch.tensor([cnt, np.uint64(fp).astype(np.int64), self.n_local, np.uint64(self.input_fingerprint).astype(np.int64)],
                           dtype=torch.int64, device=dev)
        if self.world > 1:
            dist.all_reduce(tot, group=self.group)
        t = tot.cpu().numpy()
        res = {
            "misordered": bad, "count_out_local": cnt, "eof_ok": eof_ok, "range_ok": range_ok, "status": status,
            "count_match": int(t[0]) == int(t[2]), "fingerprint_match": int(t[1]) == int(t[3]),
        }
        res["ok"] = bad == 0 and eof_ok and range_ok and status == 0 and res["count_match"] and res["fingerprint_match"]
        return res

    # ------------------------------------------------------------------------------------ failure detection
    def raise_on_failure(self) -> None:
        """Synchronises and raises :class:`~uda_b200.runtime.failure.ShuffleTimeout` when a wait or barrier kernel
        of an earlier step gave up (lost or short push, dead peer)."""
        from ..runtime.failure import check_wait_status

        torch.cuda.synchronize(self.device)
        check_wait_status(self.arena.status, f"terasort rank {self.rank}")

    def landed_segments(self):
        """``(tensor at segment start, rawLength)`` of every run of every sub-partition landed in this rank's arena."""
        if self.device_cuts and self.subparts > 1:
            raise RuntimeError("sub-segments cut on the device are slices of their parent segment: they carry no checksum of their own")
        data, segs = self.arena.data(), []
        for j in range(self.subparts):
            lay = self.plan.layout(self.rank, j)
            for s in range(self.world):
                for m in range(self.maps):
                    rid = self.plan.run_id(s, m)
                    segs.append((data[int(lay.run_offsets[rid]):], int(self.plan.index[s, m, self.rank * self.subparts + j, 1])))
        return segs

    def verify_crc(self) -> None:
        """Recomputes the IFile CRC32 of every landed segment on the device; raises ChecksumMismatch."""
        from ..utils.faults import verify_landed_crc

        verify_landed_crc(self.landed_segments())

    def close(self) -> None:
        if hasattr(self, "arena"):
            self.arena.close()
