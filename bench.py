#!/usr/bin/env python
"""bench.py — Mpackets/s of batched SFC flow lookup (BASELINE.json metric) on 1..8 B200.

Headline (`value`): configs[1] of BASELINE.json — 1024 exact-match 5-tuple rules, one 1,048,576-packet batch
of 64-byte header records per step, resident in HBM; batches rotate through a pool larger than the 126 MB L2
so every step streams from HBM.  The other workloads BASELINE.json names are measured in the same run and
reported under `configs`: configs[2] (1 Mi-flow table, 4096-packet batches, 4-NF chain), configs[3] (LPM
v4/v6 + ACL on 100 K prefixes) and configs[4] (16 Mi-flow table built on rank 0, NCCL-broadcast, 2^28 packets
sharded over the ranks, hit counters all-reduced).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--configs 1,2,3,4]

N > 1 is launched by torch.distributed.run (one rank per GPU, NCCL): rank 0 builds every table image, it is
broadcast to the other ranks, each rank classifies its own packets (no data-path collective: packets are
independent — weak scaling for the headline), and the hit counters are all-reduced.  Every rank compares one of
its own batches with the CPU oracle, and the all-reduced counters are compared with the sum of the per-rank
oracle counters; a mismatch aborts the run.
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

BATCH = 1 << 20          # packets per step (configs[1]: "1M-packet synthetic batch")
N_RULES = 1024           # configs[1]: "1K exact-match 5-tuple rules"
POOL = 8                 # distinct batches resident in HBM: 8 x 77.6 MB = 621 MB > 126 MB L2
BYTES_PER_PKT = 74       # 64 header + 2 in_port + 8 result
METRIC = "Mpackets/s of batched SFC flow lookup"
WORKLOAD = "configs[1]: 1K exact 5-tuple rules, 1M-packet batch of 64B headers, 1 launch per step"

# the gen-spec keyword sets of the four workloads; both arms build their inputs from these
CFG1_SPEC = dict(n_flows=N_RULES, n_ports=8, n_out_ports=8, hit_permille=900, nonip_permille=0, v6_permille=0,
                 vlan_permille=0, zipf=0, n_macs=8)


def cfg2_spec_kw(n_flows, n_ports):
    return dict(n_flows=n_flows, n_ports=n_ports, n_out_ports=n_ports, hit_permille=947, nonip_permille=100,
                v6_permille=200, vlan_permille=250, zipf=1, n_macs=12)


CFG3_SPEC = dict(n_flows=0, n_ports=4, port_base=32, n_out_ports=16, hit_permille=0, nonip_permille=50, v6_permille=263,
                 vlan_permille=100, n_macs=8, n_pfx4=80000, n_pfx6=20000, lpm_permille=900, n_acl=1024)


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, burst copy)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def committed_traffic(key="classify_kernel_dram_bytes_per_launch"):
    """Per-launch DRAM bytes of the classify kernel from the committed ncu capture of that workload, if any."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(key)
    except Exception:
        return None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the GPU is under the bench load."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.lines, self.proc = [], None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i", str(index)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append((time.perf_counter(), line.strip()))

    def stop(self, t0: float, t1: float):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for t, line in self.lines:
            if t < t0 or t > t1 + 0.15:
                continue
            f = [x.strip() for x in line.split(",")]
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except Exception:
                continue
            for name, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


_REAL_STDOUT = None


def quiet_stdout():
    """The contract is ONE JSON line on stdout.  Libraries under us write there too (NCCL prints its version
    banner to fd 1 on some hosts), so everything but that line is sent to stderr: fd 1 is pointed at fd 2 for the
    whole run and the line goes to a saved copy of the original descriptor."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)


def emit(obj):
    line = (json.dumps(obj) + "\n").encode()
    sys.stdout.flush()
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, line)


def log(*a):
    print("[bench]", *a, file=sys.stderr, flush=True)


# ------------------------------------------------------------------ rule sets, staged into the library and/or the oracle
class Stager:
    """Stages one rule set into the CUDA library (rank 0: `dp`, flows drawn from the product's generator `gen_dp`)
    and into the CPU oracle (every rank, rules drawn from the oracle's own generator).  Ids are handed out by the
    library in add order from zero; ranks without a handle count the same way, and rank 0 checks that the two agree."""

    def __init__(self, dp, orc_mod, gen_dp=None):
        self.dp, self.orc, self.gen, self.gen_dp = dp, orc_mod.Oracle(), orc_mod.OracleGen, gen_dp
        self.n_ports = self.n_l2 = self.n_lpm = self.n_acl = self.n_flows = 0

    def port(self, name):
        pid = self.n_ports
        self.n_ports += 1
        if self.dp is not None:
            assert self.dp.port_add(name) == pid
        return pid

    def l2(self, in_port, out_port, mac=None):
        rid = self.n_l2
        self.n_l2 += 1
        if self.dp is not None:
            assert self.dp.l2_rule_add(in_port, out_port, mac) == rid
        prio = 10 if mac is None else (100 if in_port == out_port else 32768)  # ovs-dp/ovsdp.go:120,127,135
        self.orc.l2_add(in_port, out_port, mac, prio, rid)

    def chain(self, n_vfs, n_nfs):
        """The Marvell VSP's steering set (marvell/main.go:382-403, 519-560) for a chain of n_nfs functions."""
        wire = self.port("rpm0")
        vfs = [self.port(f"vf0-{i}") for i in range(n_vfs)]
        nf_in = [self.port(f"nf{i}-in") for i in range(n_nfs)]
        nf_out = [self.port(f"nf{i}-out") for i in range(n_nfs)]
        if self.dp is not None:
            self.dp.chain_set(nf_in, nf_out)
        self.orc.chain_set(nf_in, nf_out)
        for v, port in enumerate(vfs):
            mac = self.gen.mac(v)
            self.l2(port, nf_in[0])
            self.l2(nf_in[0], port, mac)
            self.l2(nf_out[-1], nf_out[-1], mac)
            for i in range(1, n_nfs):
                self.l2(nf_in[i], nf_out[i - 1], mac)
        for i in range(n_nfs - 1):
            self.l2(nf_out[i], nf_in[i + 1])
        self.l2(nf_out[-1], wire)
        self.l2(wire, nf_out[-1])
        return self.n_ports

    def flows(self, spec, n, chunk=1 << 20):
        for first in range(0, n, chunk):
            m = min(chunk, n - first)
            keys, acts = self.gen.flows(spec, first, m)
            if self.dp is not None:
                assert self.dp.flow_add_bulk(*self.gen_dp.gen_flows(spec, first, m)) == self.n_flows
            self.orc.flow_add(keys, acts, self.n_flows)
            self.n_flows += m

    def lpm_acl(self, spec):
        for af, n in ((4, int(spec["n_pfx4"][0])), (6, int(spec["n_pfx6"][0]))):
            seen = set()
            for p, ln, port in self.gen.prefixes(spec, af, n):
                if (p, ln) in seen:  # a prefix drawn twice is staged once
                    continue
                seen.add((p, ln))
                pid = self.n_lpm
                self.n_lpm += 1
                if self.dp is not None:
                    assert self.dp.lpm_add(af, p, ln, port) == pid
                self.orc.lpm_add(af, p, ln, port, pid)
        for r in self.gen.acl_rules(spec, int(spec["n_acl"][0])):
            rid = self.n_acl
            self.n_acl += 1
            if self.dp is not None:
                assert self.dp.acl_add(r) == rid
            self.orc.acl_add(r, rid)


def run_reference(args, rank, world):
    """--impl reference: the reference has no implementation of the lookup (OVS / DPU hardware does it) and its
    rule generators are Go, which this image cannot compile; per the task's tier rules this arm times the CPU
    oracle port on all host threads of the box, on the same workload.  Nothing of the product is loaded: rules
    and packets come from the oracle's own writing of the workload (oracle/workload.cpp)."""
    if rank != 0:
        return
    import orc  # oracle bindings; never imports the product package

    st = Stager(None, orc)
    spec = orc.make_gen_spec(**CFG1_SPEC)
    st.flows(spec, N_RULES)
    threads = os.cpu_count() or 1
    pool = [orc.OracleGen.packets(spec, i * BATCH, BATCH) for i in range(min(POOL, 4))]
    for i in range(args.warmup):
        st.orc.classify(*pool[i % len(pool)], threads=threads)
    t0 = time.perf_counter()
    for i in range(args.steps):
        st.orc.classify(*pool[i % len(pool)], threads=threads)
    dt = time.perf_counter() - t0
    mpps = args.steps * BATCH / dt / 1e6
    emit({
        "impl": "reference", "metric": METRIC, "value": mpps, "unit": "Mpps", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "rules": N_RULES, "batch_packets": BATCH},
        "cpu_baseline": {"value": mpps, "unit": "Mpps", "cores": threads, "kind": "port",
                         "sample": f"{args.steps} steps x {BATCH} packets (full batch), oracle/sfc_oracle.cpp, {threads} threads"},
        "e2e": {"value": mpps, "unit": "Mpps", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference = a plain CPU port (oracle/), not a tuned CPU classifier: the Go reference has no software "
                "datapath and no Go toolchain exists here; rules and packets from oracle/workload.cpp, no product code loaded",
    })


# ------------------------------------------------------------------ the GPU arm
class Ctx:
    """What every measurement needs: ranks, modules, the stream, and the timing / parity helpers."""

    def __init__(self, args):
        import numpy as np
        import torch
        import torch.distributed as dist

        self.args, self.np, self.torch, self.dist = args, np, torch, dist
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        self.sfc = importlib.import_module("dpu-operator_b200")
        self.group = importlib.import_module("dpu-operator_b200.group")
        import orc  # the CPU oracle: the checker and the cpu_baseline, never the thing measured

        self.orc = orc
        assert torch.cuda.is_available(), "bench.py needs a B200; there is no CPU fallback"
        torch.cuda.set_device(self.local)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
        self.st = torch.cuda.current_stream().cuda_stream
        self.cpu_threads = max(1, (os.cpu_count() or 1) // self.world)
        self.peak, self.peak_src = measured_peak()

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()

    def max_over_ranks(self, v: float) -> float:
        if self.world == 1:
            return v
        t = self.torch.tensor([v], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def all_true(self, ok: bool) -> bool:
        if self.world == 1:
            return ok
        t = self.torch.tensor([1 if ok else 0], dtype=self.torch.int64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MIN)
        return bool(t.item())

    def sum_over_ranks(self, a):
        """Sum a u64 numpy array over the ranks (NCCL all-reduce of its bits as int64)."""
        if self.world == 1:
            return a
        t = self.torch.from_numpy(a.view(self.np.int64).copy()).cuda()
        self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return t.cpu().numpy().view(self.np.uint64)

    def timed(self, step, steps: int, warmup: int, warm_seconds: float = 0.0, counter=None):
        """W untimed warm-up steps (and at least warm_seconds of load, so clocks have ramped), then exactly
        `steps` steps between CUDA events, bracketed by barrier + synchronize; max over ranks.  Returns
        (milliseconds, t0, t1, launches): the host clock readings that bracket the region (for the clock sampler)
        and how far `counter()` — the library's launch count — moved inside it."""
        torch = self.torch
        t_w, i = time.perf_counter(), 0
        while i < warmup or time.perf_counter() - t_w < warm_seconds:
            step(i)
            i += 1
            if i % 64 == 0:
                torch.cuda.synchronize()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        self.barrier()
        torch.cuda.synchronize()
        c0 = counter() if counter else 0
        t0 = time.perf_counter()
        e0.record()
        for i in range(steps):
            step(i)
        e1.record()
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        c1 = counter() if counter else 0
        self.barrier()
        return self.max_over_ranks(e0.elapsed_time(e1)), t0, t1, c1 - c0

    def broadcast(self, dp):
        """Rank 0's committed image to every rank.  The first broadcast of a process also sets the NCCL
        communicator up, so the image is sent twice and the second time is the one reported."""
        if self.world == 1:
            return None
        first = self.group.broadcast_table(dp, self.rank, self.world)
        second = self.group.broadcast_table(dp, self.rank, self.world)
        return {"ms": second, "first_call_ms_incl_communicator_setup": first}

    def parity(self, name, dp, stager, spec, first_packet, n, d_hdr, d_ports, d_res, classify=None):
        """Every rank: classify `n` of its own packets (already generated on the device from `first_packet` on)
        with counters on, and compare (a) each result word with the oracle's answer on the ORACLE's own writing
        of the same packets, (b) the all-reduced hit counters with the sum of the per-rank oracle counters.
        d_hdr / d_ports are device addresses, d_res a tensor.  Aborts the run on a mismatch."""
        np, torch, sfc = self.np, self.torch, self.sfc
        dp.counters_reset()
        if classify is None:
            dp.classify(d_hdr, d_ports, d_res.data_ptr(), n, sfc.consts.F_COUNTERS, self.st)
        else:
            classify(sfc.consts.F_COUNTERS)
        torch.cuda.synchronize()
        got = d_res[:n].cpu().numpy().view(np.uint64)
        oh, op = self.orc.OracleGen.packets(spec, first_packet, n)
        want, cnt = stager.orc.classify(oh, op, threads=self.cpu_threads, counters=True)
        bad = int((got != want).sum())
        if self.world > 1 and not getattr(self, "_reduce_warm", False):
            # the first all-reduce of a process sets NCCL's channels up: do it on a scratch tensor, so that the one
            # timed below is the steady-state cost of summing the counters
            self.sum_over_ranks(np.zeros(self.sfc.consts.CNT_TOTAL, dtype=np.uint64))
            self._reduce_warm = True
        reduce_ms = self.group.reduce_counters(dp, self.world) if self.world > 1 else 0.0
        have = dp.counters_read()
        total = self.sum_over_ranks(cnt)
        counters_ok = bool(np.array_equal(have, total))
        ok = self.all_true(bad == 0 and counters_ok)
        if not ok:
            log(f"rank {self.rank}: {name}: {bad} of {n} results differ from the oracle; counters equal: {counters_ok}")
            if self.rank == 0:
                emit({"metric": METRIC, "error": f"{name}: GPU results differ from the CPU oracle", "bit_exact_all_ranks": False})
            sys.exit(1)
        return {"bit_exact_all_ranks": True, "packets_checked_per_rank": n, "counters_equal_after_allreduce": True,
                "counter_allreduce_ms": reduce_ms, "counters": {"to_cp": int(have[0]), "forward": int(have[1]), "drop": int(have[2])}}

    def roofline(self, packets_per_launch, ms_per_launch, traffic_key):
        achieved = BYTES_PER_PKT * packets_per_launch / (ms_per_launch * 1e-3) / 1e9
        return {"bound": "hbm", "achieved": achieved, "peak": self.peak, "unit": "GB/s", "frac": achieved / self.peak,
                "traffic": committed_traffic(traffic_key), "peak_source": self.peak_src,
                "algorithmic_bytes_per_launch": BYTES_PER_PKT * packets_per_launch, "kernel": "sfc::classify_kernel",
                "avg_launch_ms": ms_per_launch}


def headline(cx: Ctx):
    """configs[1]: one classify launch per 1 Mi-packet batch, 1024 exact 5-tuple rules."""
    args, sfc, torch = cx.args, cx.sfc, cx.torch
    dp = sfc.Dataplane(cx.local)
    spec = sfc.make_gen_spec(**CFG1_SPEC)
    stg = Stager(dp if cx.rank == 0 else None, cx.orc, sfc)
    stg.flows(spec, N_RULES)
    if cx.rank == 0:
        dp.commit()
    bcast = cx.broadcast(dp)
    hdr = torch.empty((POOL, BATCH, 64), dtype=torch.uint8, device="cuda")
    ports = torch.empty((POOL, BATCH), dtype=torch.int16, device="cuda")
    res = torch.zeros((POOL, BATCH), dtype=torch.int64, device="cuda")
    first = cx.rank * POOL * BATCH  # rank r streams its own slice of the synthetic packet stream
    ptrs = [(hdr[b].data_ptr(), ports[b].data_ptr(), res[b].data_ptr()) for b in range(POOL)]
    for b in range(POOL):
        dp.gen_packets(spec, first + b * BATCH, BATCH, ptrs[b][0], ptrs[b][1], cx.st)
    torch.cuda.synchronize()
    OVL = sfc.consts.F_OVERLAP

    # consecutive steps use different buffers of the pool, so they may overlap (SFC_F_OVERLAP):
    # the next launch fills SMs as the previous one drains, stream order to later ops is kept
    def step(i, flags=0):
        h, p, r = ptrs[i % POOL]
        dp.classify(h, p, r, BATCH, flags | OVL, cx.st)

    # the sampler is started before the warm-up so that the timed region follows the warm-up without a pause
    sampler = ClockSampler(cx.local) if cx.rank == 0 and not args.no_clocks else None
    ms, t0, t1, launches = cx.timed(step, args.steps, args.warmup, args.warm_seconds, lambda: dp.launch_count)
    # if the timed region was too short for nvidia-smi to see it, keep the identical load running
    # (untimed) so the clock record describes the same state
    extended = False
    if sampler is not None and (t1 - t0) < 1.0:
        extended = True
        t_end, j = time.perf_counter() + 1.2, 0
        while time.perf_counter() < t_end:
            step(j)
            j += 1
            if j % 64 == 0:
                torch.cuda.synchronize()
        torch.cuda.synchronize()
        t1 = time.perf_counter()
    clocks = sampler.stop(t0, t1) if sampler else None
    if clocks is not None:
        clocks["sampled_over"] = "timed region + 1.2 s of the identical load (region shorter than 1 s)" if extended else "timed region"
    mpps = args.steps * BATCH * cx.world / (ms * 1e-3) / 1e6

    par = cx.parity("configs[1]", dp, stg, spec, first, BATCH, ptrs[0][0], ptrs[0][1], res[0])

    # ---- e2e: the host-buffer call a dataplane shim makes (H2D + kernel + D2H inside the region) ----
    e2e = None
    if not args.no_e2e:
        n_host = 3
        h_hdr = [hdr[b].cpu().pin_memory() for b in range(n_host)]
        h_ports = [ports[b].cpu().pin_memory() for b in range(n_host)]
        h_res = [torch.zeros(BATCH, dtype=torch.int64).pin_memory() for _ in range(n_host)]
        e2e_steps = max(3, min(args.steps, 30))
        hp = [(h_hdr[b].data_ptr(), h_ports[b].data_ptr(), h_res[b].data_ptr()) for b in range(n_host)]
        for i in range(2):
            dp.classify_host_ptr(*hp[i % n_host], BATCH)

        def timed_loop(body):
            cx.barrier()
            te = time.perf_counter()
            body()
            return cx.max_over_ranks(time.perf_counter() - te)

        def one_call_per_step():  # sfc_classify_host: returns with the results in host memory
            for i in range(e2e_steps):
                dp.classify_host_ptr(*hp[i % n_host], BATCH)

        seen = []

        def queued_one_ahead():  # the loop of a dataplane: submit batch i, then collect batch i-1
            prev = None
            for i in range(e2e_steps):
                ticket = dp.classify_host_submit(*hp[i % n_host], BATCH)
                if prev is not None:
                    dp.classify_host_wait(prev[0])
                    seen.append(int(h_res[prev[1]][0]))  # the step's result is read on the host
                prev = (ticket, i % n_host)
            dp.classify_host_wait(prev[0])
            seen.append(int(h_res[prev[1]][0]))

        sync_s = timed_loop(one_call_per_step)
        e2e_s = timed_loop(queued_one_ahead)
        assert len(seen) == e2e_steps
        dp.classify(*ptrs[0], BATCH, 0, cx.st)
        torch.cuda.synchronize()
        assert torch.equal(h_res[0], res[0].cpu()), "host-buffer path and device path disagree"
        e2e = {"value": e2e_steps * BATCH * cx.world / e2e_s / 1e6, "unit": "Mpps", "h2d_bytes_per_step": BATCH * 66,
               "d2h_bytes_per_step": BATCH * 8, "steps": e2e_steps,
               "api": "sfc_classify_host_submit / _wait, batch i+1 queued before batch i is collected (pinned host buffers, "
                      "chunked H2D/kernel/D2H pipeline); every step's H2D and D2H are inside the timed region",
               "one_blocking_call_per_step_mpps": e2e_steps * BATCH * cx.world / sync_s / 1e6,
               "bound": "PCIe: 66 B in + 8 B out per packet"}

    kernel_ms = ms / args.steps  # one classify launch per step; includes launch gaps inside the region
    out = {
        "metric": METRIC, "value": mpps, "unit": "Mpps", "n_gpus": cx.world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": kernel_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "rules": N_RULES, "batch_packets": BATCH,
                   "parallelism": f"packet-sharded x{cx.world}, table replicated",
                   "submission": "one launch per batch, back-to-back launches overlap their ramp-up/drain (programmatic dependent launch)",
                   "l2": f"inputs rotate through {POOL} distinct batches ({POOL * BATCH * BYTES_PER_PKT / 1e6:.0f} MB) > 126 MB L2"},
        "roofline": cx.roofline(BATCH, kernel_ms, "configs1_dram_bytes_per_launch"),
        "clocks": clocks, "gpu_launches": launches,
        "bit_exact_all_ranks": par["bit_exact_all_ranks"], "parity": par,
        "table": {"bytes": dp.table_size, "broadcast": bcast},
    }
    if e2e:
        out["e2e"] = e2e
    # ---- CPU baseline: the oracle port on the box's host cores, bounded sample ----
    if cx.world == 1 and not args.no_cpu_baseline:
        oh, op = cx.orc.OracleGen.packets(spec, first, BATCH)
        res_cpu = {}
        for threads in sorted({1, os.cpu_count() or 1}):
            t_start, passes, best = time.perf_counter(), 0, 1e9
            while time.perf_counter() - t_start < (4.0 if threads == 1 else 6.0):
                ta = time.perf_counter()
                stg.orc.classify(oh, op, threads=threads)
                best = min(best, time.perf_counter() - ta)
                passes += 1
            res_cpu[threads] = (BATCH / best / 1e6, passes)
        nthr = max(res_cpu)
        out["cpu_baseline"] = {"value": res_cpu[nthr][0], "unit": "Mpps", "cores": nthr, "kind": "port",
                               "sample": f"the same 1M-packet batch, best of {res_cpu[nthr][1]} passes (~6 s), oracle/sfc_oracle.cpp "
                                         "(a plain port, not a tuned CPU classifier)",
                               "single_thread_mpps": res_cpu[1][0], "bit_exact_vs_gpu": True}
    del hdr, ports, res, dp
    torch.cuda.empty_cache()
    return out


# ------------------------------------------------------------------ the other workloads of BASELINE.json
def _steps(cx):
    """Steps of a `configs` record: these are bounded side measurements, not the K of the headline — long enough
    (a few milliseconds) that the fill and drain of the launch train do not dominate."""
    return 200 if cx.args.steps >= 20 else max(4, cx.args.steps)


def record_cfg2(cx: Ctx):
    """configs[2]: 1 Mi-flow table (IPv4 + IPv6, Zipf), 4-NF chain on 8 VFs, 4096-packet batches; one
    sfc_classify_batches call carries 256 separate batches (1 Mi packets) and is one step."""
    sfc, torch, np = cx.sfc, cx.torch, cx.np
    NB, B, REGIONS, NF = 256, 4096, 4, 1 << 20
    dp = sfc.Dataplane(cx.local)
    stg = Stager(dp if cx.rank == 0 else None, cx.orc, sfc)
    n_ports = stg.chain(8, 4)
    spec = sfc.make_gen_spec(**cfg2_spec_kw(NF, n_ports))
    t_b = time.perf_counter()
    stg.flows(spec, NF)
    if cx.rank == 0:
        dp.commit()
    build_s = time.perf_counter() - t_b
    bcast = cx.broadcast(dp)
    n = NB * B
    hdr = torch.empty((REGIONS, n, 64), dtype=torch.uint8, device="cuda")
    ports = torch.empty((REGIONS, n), dtype=torch.int16, device="cuda")
    res = torch.zeros((REGIONS, n), dtype=torch.int64, device="cuda")
    first = cx.rank * REGIONS * n
    descs = []
    for r in range(REGIONS):
        dp.gen_packets(spec, first + r * n, n, hdr[r].data_ptr(), ports[r].data_ptr(), cx.st)
        d = np.zeros(NB, dtype=sfc.BATCH_DTYPE)
        hp, pp, rp = hdr[r].data_ptr(), ports[r].data_ptr(), res[r].data_ptr()
        for i in range(NB):
            d[i] = (hp + i * B * 64, pp + i * B * 2, rp + i * B * 8, B, 0)
        descs.append(d)
    torch.cuda.synchronize()
    OVL = sfc.consts.F_OVERLAP
    steps = _steps(cx)
    ms, _, _, launches = cx.timed(lambda i: dp.classify_batches(descs[i % REGIONS], OVL, cx.st), steps, 5, 0.0, lambda: dp.launch_count)
    # the same packets as ONE contiguous launch per step, for comparison
    ptrs = [(hdr[r].data_ptr(), ports[r].data_ptr(), res[r].data_ptr()) for r in range(REGIONS)]
    ms1, _, _, _ = cx.timed(lambda i: dp.classify(*ptrs[i % REGIONS], n, OVL, cx.st), steps, 5)
    par = cx.parity("configs[2]", dp, stg, spec, first, n, None, None, res[0],
                    classify=lambda flags: dp.classify_batches(descs[0], flags, cx.st))
    rec = {"workload": "configs[2]: 1 Mi-flow table (20 % IPv6, Zipf, 5 % misses, 10 % non-IP, 25 % tagged), 4-NF chain on 8 VFs, "
                       "256 separate 4096-packet batches per sfc_classify_batches call",
           "mpps": steps * n * cx.world / (ms * 1e-3) / 1e6, "steps": steps, "ms_per_step": ms / steps, "gpu_launches": launches,
           "roofline": cx.roofline(n, ms / steps, "configs2_dram_bytes_per_launch"),
           "one_contiguous_launch_mpps": steps * n * cx.world / (ms1 * 1e-3) / 1e6,
           "table": {"bytes": dp.table_size, "flows": NF, "build_seconds": build_s, "broadcast": bcast}, **par}
    del hdr, ports, res, dp, stg
    torch.cuda.empty_cache()
    return rec


def record_cfg3(cx: Ctx):
    """configs[3]: LPM (80 K IPv4 + 20 K IPv6 prefixes) + a 1024-rule ACL, 70 / 25 / 5 % IPv4 / IPv6 / non-IP."""
    sfc, torch = cx.sfc, cx.torch
    REGIONS, n = 4, 1 << 20
    dp = sfc.Dataplane(cx.local)
    stg = Stager(dp if cx.rank == 0 else None, cx.orc, sfc)
    spec = sfc.make_gen_spec(**CFG3_SPEC)
    t_b = time.perf_counter()
    stg.lpm_acl(spec)
    if cx.rank == 0:
        dp.commit()
    build_s = time.perf_counter() - t_b
    bcast = cx.broadcast(dp)
    hdr = torch.empty((REGIONS, n, 64), dtype=torch.uint8, device="cuda")
    ports = torch.empty((REGIONS, n), dtype=torch.int16, device="cuda")
    res = torch.zeros((REGIONS, n), dtype=torch.int64, device="cuda")
    first = cx.rank * REGIONS * n
    for r in range(REGIONS):
        dp.gen_packets(spec, first + r * n, n, hdr[r].data_ptr(), ports[r].data_ptr(), cx.st)
    torch.cuda.synchronize()
    OVL = sfc.consts.F_OVERLAP
    steps = _steps(cx)
    ptrs = [(hdr[r].data_ptr(), ports[r].data_ptr(), res[r].data_ptr()) for r in range(REGIONS)]
    ms, _, _, launches = cx.timed(lambda i: dp.classify(*ptrs[i % REGIONS], n, OVL, cx.st), steps, 5, 0.0, lambda: dp.launch_count)
    m = 1 << 17  # the oracle scans the 1024 ACL rules linearly: a bounded sample per rank
    par = cx.parity("configs[3]", dp, stg, spec, first, m, ptrs[0][0], ptrs[0][1], res[0])
    rec = {"workload": "configs[3]: LPM on 80 K IPv4 + 20 K IPv6 prefixes + 1024-rule ACL (5-field ternary, 10 % deny), "
                       "70 / 25 / 5 % IPv4 / IPv6 / non-IP, one launch per 1 Mi packets",
           "mpps": steps * n * cx.world / (ms * 1e-3) / 1e6, "steps": steps, "ms_per_step": ms / steps, "gpu_launches": launches,
           "roofline": cx.roofline(n, ms / steps, "configs3_dram_bytes_per_launch"),
           "table": {"bytes": dp.table_size, "prefixes": stg.n_lpm, "acl_rules": stg.n_acl, "build_seconds": build_s, "broadcast": bcast}, **par}
    del hdr, ports, res, dp, stg
    torch.cuda.empty_cache()
    return rec


def record_cfg4(cx: Ctx):
    """configs[4]: a 16 Mi-flow table built on rank 0 and broadcast, 2^28 packets sharded
    contiguously over the ranks (strong scaling), per-action / per-stage / per-rule counters on, all-reduced."""
    sfc, torch = cx.sfc, cx.torch
    NF, TOTAL, M = 1 << 24, 1 << 28, 1 << 20
    dp = sfc.Dataplane(cx.local)
    stg = Stager(dp if cx.rank == 0 else None, cx.orc, sfc)
    n_ports = stg.chain(8, 4)
    spec = sfc.make_gen_spec(**cfg2_spec_kw(NF, n_ports))
    t_b = time.perf_counter()
    stg.flows(spec, NF)
    if cx.rank == 0:
        dp.commit()
    build_s = time.perf_counter() - t_b
    bcast = cx.broadcast(dp)
    lo, hi = cx.group.shard_range(TOTAL, cx.rank, cx.world)
    n = hi - lo
    hdr = torch.empty((n, 64), dtype=torch.uint8, device="cuda")
    ports = torch.empty(n, dtype=torch.int16, device="cuda")
    res = torch.zeros(n, dtype=torch.int64, device="cuda")
    hp, pp, rp = hdr.data_ptr(), ports.data_ptr(), res.data_ptr()
    for o in range(0, n, 1 << 22):
        m = min(1 << 22, n - o)
        dp.gen_packets(spec, lo + o, m, hp + o * 64, pp + o * 2, cx.st)
    torch.cuda.synchronize()
    flags = sfc.consts.F_OVERLAP | sfc.consts.F_COUNTERS
    nl = n // M

    def whole_shard(_i):
        for k in range(nl):
            dp.classify(hp + k * M * 64, pp + k * M * 2, rp + k * M * 8, M, flags, cx.st)

    dp.counters_reset()
    ms, _, _, _ = cx.timed(whole_shard, 1, 1)  # one pass over the rank's whole shard is the timed unit (after one untimed pass)
    cnt = dp.counters_read(16)
    counted = int(cnt[0] + cnt[1] + cnt[2])
    assert counted == 2 * n, f"counters saw {counted} packets, the two passes hold {2 * n}"
    par = cx.parity("configs[4]", dp, stg, spec, lo, M, hp, pp, res)
    rec = {"workload": "configs[4]: 16 Mi-flow table (20 % IPv6) built on rank 0 and broadcast, Zipf traffic, 2^28 packets "
                       "sharded contiguously over the ranks, counters on, one launch per 1 Mi packets",
           "scaling": "strong", "mpps": TOTAL / (ms * 1e-3) / 1e6, "packets": TOTAL, "packets_per_rank": n, "ms_total": ms,
           "gpu_launches": nl,
           "roofline": dict(cx.roofline(M, ms / nl, "configs4_dram_bytes_per_launch"), note="per GPU"),
           "table": {"bytes": dp.table_size, "flows": NF, "build_seconds": build_s, "broadcast": bcast}, **par}
    del hdr, ports, res, dp, stg
    torch.cuda.empty_cache()
    return rec


def main():
    quiet_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--configs", default="1,2,3,4", help="which BASELINE.json configs to measure (1 is the headline and always runs)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--warm-seconds", type=float, default=0.3, help="minimum load time before the timed region (clock ramp)")
    ap.add_argument("--no-clocks", action="store_true", help="skip nvidia-smi sampling (profiler runs)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args, int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")))
    cx = Ctx(args)
    out = headline(cx)
    want = {c.strip() for c in args.configs.split(",")}
    records = {}
    for key, fn in (("2", record_cfg2), ("3", record_cfg3), ("4", record_cfg4)):
        if key in want:
            t = time.perf_counter()
            records[f"configs[{key}]"] = fn(cx)
            log(f"rank {cx.rank}: configs[{key}] measured in {time.perf_counter() - t:.1f} s")
    if cx.rank == 0:
        out["configs"] = records
        out["bit_exact_all_ranks"] = all([out["bit_exact_all_ranks"]] + [r["bit_exact_all_ranks"] for r in records.values()])
        emit(out)
    if cx.world > 1:
        cx.dist.destroy_process_group()


if __name__ == "__main__":
    main()
