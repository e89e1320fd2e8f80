#!/usr/bin/env python
"""bench.py -- M17 baseband Msamples/s demodulated on N B200 (driver contract).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c3|c2] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A step = every channel of this rank advances by `chunk` samples (one m17b_submit of
[channels][chunk] int16 + one drain of the decoded frame records).  Channels are continuous
across steps (state stays resident on the GPU).  `value` times K steps with the inputs already
in HBM; `e2e` times the same call sequence from pinned host memory, host->device copies and the
device->host read of the frame records included.  Streams shard across ranks with no data-path
collective (weak scaling: `channels` per GPU); only counters are gathered.

--impl reference times the reference's own CPU implementation (oracle/_ref/m17-demod-ref, the
unmodified reference headers, one forked process per stream over all host cores) on a bounded
sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ALG_FLOP_PER_SAMPLE = 300.0          # 150 multiplies + 150 adds per filtered sample


def kernel_source_hash():
    """sha256 over the kernel sources: what a committed ncu capture is stamped with (the GPU box has no .git)."""
    import hashlib
    h = hashlib.sha256()
    for f in ("m17b_kernels.cu", "m17b_lane.cuh", "m17b_dev.cuh"):
        h.update(open(os.path.join(ROOT, "m17_cxx_demod_b200", "csrc", f), "rb").read())
    return h.hexdigest()[:16]


def fir_traffic(channels, chunk, fma):
    """DRAM bytes per fir_kernel launch from the committed ncu --set full capture (profiles/fir_dram_traffic.json), or
    None when that capture was taken on another launch shape or on other kernel sources than the ones that run here."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "fir_dram_traffic.json")))
    except OSError:
        return None
    if fma or t["channels"] != channels or t["chunk_samples"] != chunk or t.get("kernel_source_hash") != kernel_source_hash():
        return None
    return t["dram_bytes_read"] + t["dram_bytes_write"]


def bind_to_gpu_numa_node(index):
    """Run this rank (and first-touch its pinned staging buffers) on the CPUs NVML reports as local to its GPU, so
    that N ranks uploading at once do not all cross the same inter-socket link.  No-op when NVML gives no answer."""
    try:
        import pynvml
        pynvml.nvmlInit()
        n_cpu = os.cpu_count() or 1
        mask = pynvml.nvmlDeviceGetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(index), (n_cpu + 63) // 64)
        cpus = {i for i in range(n_cpu) if (mask[i // 64] >> (i % 64)) & 1} & os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
    except Exception:
        pass


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c3", choices=["c2", "c3", "c4"])
    ap.add_argument("--channels", type=int, default=0, help="channels per GPU (default: workload's)")
    ap.add_argument("--chunk", type=int, default=9600, help="samples per channel per step")
    ap.add_argument("--unique", type=int, default=1024, help="unique waveforms tiled over the channels")
    ap.add_argument("--gen", default="cpu", choices=["cpu", "device"],
                    help="device: the waveforms are built on the GPU by the device-side modulator (bit-identical to the CPU "
                         "generator; makes --unique up to the channel count practical: every channel its own waveform)")
    ap.add_argument("--fma", action="store_true", help="fused multiply-add matched filter (not bit-identical)")
    ap.add_argument("--e2e-steps", type=int, default=8)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--parity-waveforms", type=int, default=4096, help="unique waveforms the reference checks in the parity leg")
    ap.add_argument("--no-overlap", action="store_true",
                    help="diagnostic: wait for every step's own records before the next submit (no step overlaps the next)")
    ap.add_argument("--ref-streams", type=int, default=1024, help="streams per step of the reference arm")
    return ap.parse_args()


def workload(args, total_samples):
    from m17_cxx_demod_b200 import gen
    seconds = total_samples / 48000.0
    if args.workload == "c2":
        w = gen.workload_c2(n_streams=args.unique, seconds=seconds)
        channels = args.channels or 1024
        desc = "configs[1]: stream-mode channels, AWGN Eb/N0 3-12 dB swept over channels"
    elif args.workload == "c4":
        w = gen.workload_c4(n_unique=args.unique, seconds=seconds)
        channels = args.channels or 65536
        desc = "configs[3]: mixed LSF/stream/packet/BERT traffic, SNR 15 dB"
    else:
        w = gen.workload_c3(n_unique=args.unique, seconds=seconds)
        channels = args.channels or 65536
        desc = "configs[2]: 65,536 channels, +-500 ppm clock skew, +-0.1 frequency offset, SNR 20 dB"
    w.n_samples = total_samples
    return w, channels, desc


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}", "--format=csv,noheader,nounits",
                                          "-lms", "20"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx = float(f[1])
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        busy = [v for v in sm if v > 0.5 * (mx or 1)] or sm
        return {"sm_mhz": float(np.median(busy)) if busy else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def run_reference(args):
    """The reference's CPU path on this box's host cores: rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as ol
    from m17_cxx_demod_b200 import gen
    total = (args.steps + args.warmup) * args.chunk
    w, channels, desc = workload(args, total)
    cores = os.cpu_count() or 1
    n_ref = min(args.ref_streams, len(w.params))
    if not os.path.exists(ol.ref_binary()):
        kind = "port"
    else:
        kind = "reference"
    x = gen.generate(w.params[:n_ref], total)
    times = []
    records = 0
    with tempfile.TemporaryDirectory(dir="/dev/shm" if os.path.isdir("/dev/shm") else None) as td:
        path = os.path.join(td, "x.s16")
        x.tofile(path)
        for step in range(args.warmup + args.steps):
            t0 = time.perf_counter()
            if kind == "reference":
                out = subprocess.run([ol.ref_binary(), "--batch", path, str(n_ref), str(total), "--jobs", str(cores), "--quiet"],
                                     capture_output=True, text=True, check=True)
                info = json.loads(out.stderr.strip().splitlines()[-1])
                dt = info["seconds"]; records = info["records"]
            else:
                rec, _ = ol.demod_batch("port", x)
                dt = time.perf_counter() - t0; records = sum(len(r) for r in rec)
            if step >= args.warmup:
                times.append(dt)
    ms = 1e3 * float(np.mean(times))
    value = n_ref * total / (ms * 1e-3) / 1e6
    sample = f"{n_ref} of the workload's channels x {total} samples each per step, one forked m17-demod-ref process per stream, {cores} in flight"
    line = {
        "impl": "reference", "metric": "baseband_msamples_per_s", "value": value, "unit": "Msamples/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": desc, "channels_per_gpu": channels, "chunk_samples": args.chunk, "unique_waveforms": len(w.params)},
        "frames_per_s": records / (ms * 1e-3),
        "cpu_baseline": {"value": value, "unit": "Msamples/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
        return

    import torch
    import torch.distributed as dist
    import m17_cxx_demod_b200 as m
    from m17_cxx_demod_b200 import gen
    from m17_cxx_demod_b200.shard import shard_range, gather_counters

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        bind_to_gpu_numa_node(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)

    K, W, T = args.steps, args.warmup, args.chunk
    P = 3                                                      # profiled steps after the timed ones (kernel shares)
    total = (K + W + P) * T
    w, channels, desc = workload(args, total)
    U = len(w.params)
    lo, hi = shard_range(channels * world, world, rank)       # weak scaling: `channels` per GPU
    assert hi - lo == channels

    # ---- synthetic input: U unique waveforms generated on the CPU, tiled on the device -------
    t0 = time.perf_counter()
    if args.gen == "device":
        # every waveform built on the GPU (csrc/m17gen_dev.cu); the host copy is only for the parity leg and e2e
        uniq_d = torch.empty((U, total), dtype=torch.int16, device=dev)
        gen.generate_device(w.params, total, uniq_d.data_ptr(), device=local)
        torch.cuda.synchronize()
        gen_s = time.perf_counter() - t0
        uniq = uniq_d.cpu().numpy()
    else:
        uniq = gen.generate(w.params, total)
        gen_s = time.perf_counter() - t0
        uniq_d = torch.from_numpy(uniq).to(dev)
    if U == channels and lo % U == 0:
        big = uniq_d                                           # every channel its own waveform
    else:
        idx = (torch.arange(lo, hi, device=dev) % U)
        big = uniq_d.index_select(0, idx)                      # [channels][total] int16 resident in HBM
    assert big.is_contiguous()
    torch.cuda.synchronize()

    flags = m.STREAMING | (m.FIR_FMA if args.fma else 0)      # per-kernel event records stay off in the timed steps
    rec_cap = min(channels * 2 * (T // 1600 + 2) * 3, 2**31 - 1)          # two steps in flight
    d = m.Demodulator(channels, T, flags=flags, device=local, frame_capacity=rec_cap)
    fp32_peak = d.measure_fp32_peak(1)
    mix_scalar, mix_packed = d.measure_fp32_peak(2), d.measure_fp32_peak(4)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # m17b_submit only enqueues (filter, then rounds of scan/decode/fix-up kernels).  Step k+1 is enqueued
    # before the frame records of step k are read back (m17b_drain_upto waits for step k only), so the
    # device never idles on the host; every step's records are read back inside the timed region.
    n_frames = 0
    records = []
    for k in range(W):
        d.submit_device(big.data_ptr() + 2 * k * T, T, total)
        records.append(d.drain())
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_wall0 = time.perf_counter()
    ev0.record()
    d.event_record(0)
    prev = None
    for k in range(W, W + K):
        d.submit_device(big.data_ptr() + 2 * k * T, T, total)
        t_k = d.ticket()
        if args.no_overlap:
            r = d.drain_upto(t_k)
            n_frames += len(r)
            records.append(r)
            continue
        if prev is not None:
            r = d.drain_upto(prev)
            n_frames += len(r)
            records.append(r)
        prev = t_k
    r = d.drain()
    n_frames += len(r)
    records.append(r)
    d.event_record(1)
    ev1.record()
    barrier()
    t_wall = time.perf_counter() - t_wall0
    t_dev = d.event_elapsed_ms(0, 1) * 1e-3      # CUDA events on the library's compute stream
    clocks = sampler.stop()
    ctr_timed = d.counters()
    # Kernel shares: P more steps of the same stream of samples, enqueued the same way, with the per-kernel event
    # records on (two event records around every launch: kept out of the timed steps).
    d.profile_enable(True)
    d.profile(reset=True)
    prev = None
    for k in range(W + K, W + K + P):
        d.submit_device(big.data_ptr() + 2 * k * T, T, total)
        t_k = d.ticket()
        if prev is not None:
            records.append(d.drain_upto(prev))
        prev = t_k
    records.append(d.drain())
    prof = d.profile()
    ctr = d.counters()
    # Timed on the device (events on the launching stream); the host clock around the same K steps
    # is reported beside it (m17b_drain() synchronises every step, so the two agree closely).
    t = torch.tensor([t_dev], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    t_max = float(t.item())
    totals = gather_counters({"frames": n_frames, "samples": channels * K * T}, world, dev if world > 1 else None)

    value = totals["samples"] / t_max / 1e6
    # Roofline probe.  In M17B_STREAMING mode the rounds of step k share the SMs with the matched filter
    # of step k+1, so the filter's in-step launch time includes that time-sharing.  Three more submits, each
    # followed by m17b_sync, time the same launch (same shapes, same resident data) with nothing else running.
    fir_in_step_ms = prof.fir_ms / max(1, prof.fir_launches)
    fir_samples_per_launch = prof.fir_samples / max(1, prof.fir_launches)
    d.profile(reset=True)
    for k in range(3):
        d.submit_device(big.data_ptr() + 2 * k * T, T, total)
        d.sync()
    d.drain()
    prof_alone = d.profile()
    fir_ms_per_launch = prof_alone.fir_ms / max(1, prof_alone.fir_launches)
    achieved = ALG_FLOP_PER_SAMPLE * fir_samples_per_launch / (fir_ms_per_launch * 1e-3) / 1e12 if fir_ms_per_launch else 0.0
    kernel_ms = {"fir": prof.fir_ms, "dcd": prof.dcd_ms, "scan": prof.scan_ms, "scan_acquiring": prof.scan1_ms, "decode": prof.decode_ms, "fixup_search": prof.fixup_ms,
                 "tail": prof.aux_ms}
    ksum = sum(kernel_ms.values()) or 1.0
    launches_per_step = (prof.fir_launches + prof.dcd_launches + prof.aux_launches + prof.scan_launches + prof.decode_launches +
                         prof.fixup_launches) / P
    launches = int(round(launches_per_step * K))
    # achieved HBM figures of the latency-bound kernels (north_star asks for them): algorithmic bytes per step over the
    # kernels' summed launch time (events of the profiled steps; launches of different chains overlap, so these are lower bounds)
    try:
        hbm_peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
        hbm_src = "MEASURED_PEAKS.json hbm_gbs"
    except (OSError, KeyError, ValueError):
        hbm_peak, hbm_src = 6550.0, "fallback stated in B200_PROFILING.md"
    frames_step = n_frames / K
    samples_step = channels * T

    def hbm_entry(alg_bytes, ms, n_launch, what):
        gbs = alg_bytes / (ms * 1e-3) / 1e9 if ms else None
        return {"algorithmic_bytes_per_step": alg_bytes, "what": what, "kernel_ms_per_step": ms, "launches_per_step": n_launch,
                "achieved_gbs": gbs, "frac_of_hbm_peak": gbs / hbm_peak if gbs else None, "bound": "latency (dependent issue), not HBM"}
    roofline_hbm = {
        "peak_gbs": hbm_peak, "peak_source": hbm_src,
        "dcd": hbm_entry(2.0 * samples_step, prof.dcd_ms / P, prof.dcd_launches / P, "2 B/sample read (the int16 input; x[n-120] comes from the same rows again, out of L1 / L2)"),
        "scan": hbm_entry(4.0 * samples_step + 368.0 * frames_step, (prof.scan_ms + prof.scan1_ms) / P, prof.scan_launches / P,
                          "4 B/sample read (matched-filter output) + 368 B/frame written (soft bits)"),
        "decode": hbm_entry((368.0 + 80.0) * frames_step, prof.decode_ms / P, prof.decode_launches / P,
                            "368 B/frame read (soft bits) + 80 B/record written"),
    }

    line = {
        "metric": "baseband_msamples_per_s", "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": 1e3 * t_max / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": desc, "channels_per_gpu": channels, "chunk_samples": T, "unique_waveforms": U,
                   "filter": "fma" if args.fma else "exact (reference mul-then-add order, bit-identical)",
                   "l2": f"inputs larger than L2 ({channels * T * 2 / 1e6:.0f} MB of int16 per step, never re-read)",
                   "drain": "records of step k read back (m17b_drain_upto) while step k+1 runs; all inside the timed region",
                   "mode": "M17B_STREAMING (streams walk whole frames; every frame ending in a step is delivered with that step)",
                   "generator": "device-side modulator (csrc/m17gen_dev.cu)" if args.gen == "device" else "CPU generator (csrc/m17gen.cpp)",
                   "generator_s": round(gen_s, 2), **({"no_overlap": True} if args.no_overlap else {})},
        "frames_per_s": totals["frames"] / t_max, "wall_s": t_wall, "device_s": t_dev,
        "roofline": {"bound": "fp32", "kernel": "fir_kernel", "achieved": achieved, "peak": fp32_peak, "unit": "TFLOP/s",
                     "frac": achieved / fp32_peak if fp32_peak else None, "traffic": fir_traffic(channels, T, args.fma),
                     "peak_source": "measured here: packed FFMA microkernel (m17b_measure_fp32_peak); MEASURED_PEAKS.json has no FP32 entry",
                     "alg_flop_per_sample": ALG_FLOP_PER_SAMPLE, "avg_launch_ms": fir_ms_per_launch,
                     "timing": "CUDA events around 3 fir_kernel launches run alone right after the timed region (same shapes and data)",
                     "in_step": {"avg_launch_ms": fir_in_step_ms,
                                 "frac": (achieved * fir_ms_per_launch / fir_in_step_ms / fp32_peak) if fp32_peak and fir_in_step_ms else None,
                                 "note": "inside the timed steps the launch shares the SMs with the previous step's scan / decode rounds (M17B_STREAMING runs the sample front end of step k+1 on its own stream under the rounds of step k)"},
                     # the bit-identical filter must round the product and the sum separately: a multiply and an
                     # add per tap instead of one FFMA, i.e. at most half the FFMA-counted peak.  Measured here
                     # with the same instruction mix (FMUL+FADD, and 2 FMUL + 1 packed FADD2 as the kernel issues).
                     "exact_mode_ceiling": None if args.fma else {
                         "fmul_fadd_tflops": mix_scalar, "fmul_fadd2_tflops": mix_packed,
                         "frac_of_ceiling": achieved / max(mix_scalar, mix_packed) if max(mix_scalar, mix_packed) > 0 else None}},
        "kernel_share": {k: v / ksum for k, v in kernel_ms.items()},
        "kernel_ms_per_step": {k: v / P for k, v in kernel_ms.items()},
        "kernel_timing": f"CUDA events around every launch during {P} further steps after the timed ones (none inside the timed steps); launches of the two chains and of the next step's front end overlap, so the sum exceeds the step",
        "roofline_hbm": roofline_hbm,
        "gpu_launches": launches,
        "search": {"requests_per_step": prof.search_requests / P, "samples_per_step": prof.search_samples / P,
                   "note": "streams in ST_UNLOCKED with the carrier detect on ask for a parallel sync-word search over the samples ahead"},
        "clocks": clocks,
        "frames_by_type": list(ctr_timed.frames), "faults": int(ctr.faults), "dropped": int(ctr.dropped),
    }

    # ---- parity of the benchmarked run itself: every channel against the oracle of its waveform ----
    if rank == 0 and not args.no_parity:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import oracle_lib as ol
        # every waveform, hence every channel -- up to --parity-waveforms of them: the reference demodulates about one
        # million samples per second and core, so 65,536 unique waveforms would take it ten minutes
        n_chk = min(U, args.parity_waveforms)
        want, _ = ol.demod_batch("ref" if ol.have_ref() else "port", uniq[:n_chk, :])
        allrec = np.concatenate(records) if records else np.zeros(0, m.FRAME_DTYPE)
        per = m.demod.split_by_stream(allrec, channels)
        bad = 0
        bad_rec = 0
        cmp_rec = 0
        checked = 0
        examples = []
        for s in range(channels):
            u = (lo + s) % U
            if u >= n_chk:
                continue
            checked += 1
            cmp_rec += len(per[s])
            if not ol.records_equal(want[u], per[s]):
                bad += 1
                a, b = want[u], per[s]
                n = min(len(a), len(b))
                same = np.ones(n, bool)
                for f in ("seq", "viterbi_cost", "type", "payload", "lsf", "sample_pos"):
                    same &= (a[f][:n] == b[f][:n]) if a[f].ndim == 1 else (a[f][:n] == b[f][:n]).all(axis=1)
                bad_rec += int((~same).sum()) + abs(len(a) - len(b))
                if len(examples) < 4:
                    examples.append(f"channel {s} (waveform {u}): " + ol.describe_mismatch(want[u], per[s]))
        line["parity"] = {"channels": channels, "channels_checked": checked, "mismatching": bad, "records_compared": int(cmp_rec),
                          "mismatching_records": bad_rec,
                          "oracle": "ref" if ol.have_ref() else "port",
                          "what": f"all records of all {K + W + P} steps of every checked channel against the oracle's records of its waveform, all fields bit for bit"}
        if examples:
            line["parity"]["examples"] = examples

    del d
    # ---- e2e: same call sequence from pinned host memory --------------------------------------
    if not args.no_e2e:
        Ke = max(1, args.e2e_steps)
        try:                                                   # bound the pinned host memory of all ranks to a quarter of what is free
            import psutil
            budget = psutil.virtual_memory().available // 4 // max(1, world)
            Ke = max(2, min(Ke, int(budget // (channels * T * 2)) - 2))
        except ImportError:
            Ke = Ke if world == 1 else min(Ke, 4)
        NW = 2                                                 # untimed warm-up submits (one per staging buffer)
        if (Ke + NW) * T > total:
            Ke = total // T - NW
        DEPTH = 2                                              # submits enqueued ahead of the record read-back
        de = m.Demodulator(channels, T, flags=m.STREAMING, device=local,
                           frame_capacity=min(channels * (DEPTH + 2) * (T // 1600 + 2) * 3, 2**31 - 1))
        bufs = [m.demod.PinnedBuffer((channels, T)) for _ in range(Ke + NW)]
        idx_h = (np.arange(lo, hi) % U)
        for k, b in enumerate(bufs):
            np.take(uniq[:, k * T:(k + 1) * T], idx_h, axis=0, out=b.array)
        for k in range(NW):
            de.submit(bufs[k].array)
            de.drain()
        barrier()
        t0 = time.perf_counter()
        marks = []
        d2h = 0
        fr = 0
        # Every step: one m17b_submit from pinned host memory (H2D on the copy stream into one of two
        # staging buffers, kernels on the compute stream) and one read-back of that step's frame records
        # (D2H on its own stream).  Submits are enqueued DEPTH steps ahead of the read-back, so the upload
        # of chunk k+2 runs under the kernels of chunk k+1 and the read-back of chunk k.
        tickets = []
        for k in range(NW, Ke + NW):
            de.submit(bufs[k].array)
            tickets.append(de.ticket())
            if len(tickets) > DEPTH:
                r = de.drain_upto(tickets.pop(0))
                marks.append(time.perf_counter())
                d2h += r.nbytes
                fr += len(r)
        while tickets:
            t_k = tickets.pop(0)
            r = de.drain_upto(t_k) if tickets else de.drain()
            marks.append(time.perf_counter())
            d2h += r.nbytes
            fr += len(r)
        barrier()
        te = time.perf_counter() - t0
        # upload rate and kernel times of this mode from two more submits with the event records on (outside the timing)
        de.profile_enable(True)
        de.profile(reset=True)
        for k in range(2):
            de.submit(bufs[k].array)
        de.drain()
        prof_e = de.profile()
        Kp = 2
        tt = torch.tensor([te], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        te = float(tt.item())
        line["e2e"] = {"value": world * channels * Ke * T / te / 1e6, "unit": "Msamples/s",
                       "h2d_bytes_per_step": channels * T * 2, "d2h_bytes_per_step": d2h // Ke, "steps": Ke,
                       "frames_per_s": world * fr / te, "pipeline_depth": DEPTH,
                       "ms_between_drains": [round(1e3 * (b - a), 2) for a, b in zip([t0] + marks[:-1], marks)],
                       "kernel_ms_per_step": {"fir": prof_e.fir_ms / Kp, "dcd": prof_e.dcd_ms / Kp, "scan": prof_e.scan_ms / Kp,
                                              "scan_acquiring": prof_e.scan1_ms / Kp, "decode": prof_e.decode_ms / Kp, "h2d": prof_e.h2d_ms / Kp},
                       "h2d_gbs": (prof_e.h2d_bytes / (prof_e.h2d_ms * 1e-3) / 1e9) if prof_e.h2d_ms else None,
                       "timing": "host clock around Ke submit+drain pairs incl. pipeline fill and final drain"}
        for b in bufs:
            b.free()
        del de

    # ---- CPU baseline: the reference's own code on this box's host cores (rank 0, N = 1 only) ----
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import oracle_lib as ol
        cores = os.cpu_count() or 1
        n_ref = min(U, 1024)
        with tempfile.TemporaryDirectory(dir="/dev/shm" if os.path.isdir("/dev/shm") else None) as td:
            path = os.path.join(td, "x.s16")
            uniq[:n_ref].tofile(path)
            if os.path.exists(ol.ref_binary()):
                # repeat the batch until about 10 s of CPU wall time have been measured
                secs, reps = 0.0, 0
                while secs < 10.0 and reps < 64:
                    out = subprocess.run([ol.ref_binary(), "--batch", path, str(n_ref), str(total), "--jobs", str(cores), "--quiet"],
                                         capture_output=True, text=True, check=True)
                    info = json.loads(out.stderr.strip().splitlines()[-1])
                    secs += info["seconds"]; reps += 1
                line["cpu_baseline"] = {"value": reps * n_ref * total / secs / 1e6, "unit": "Msamples/s", "cores": cores, "kind": "reference",
                                        "sample": f"{reps} x ({n_ref} of the workload's unique channels x {total} samples), one m17-demod-ref "
                                                  f"process per stream, {cores} in flight, {secs:.2f} s wall"}
            else:
                t0 = time.perf_counter()
                ol.demod_batch("port", uniq[:n_ref])
                dt = time.perf_counter() - t0
                line["cpu_baseline"] = {"value": n_ref * total / dt / 1e6, "unit": "Msamples/s", "cores": 1, "kind": "port",
                                        "sample": f"{n_ref} channels x {total} samples, scalar port, {dt:.2f} s wall"}

    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
