#!/usr/bin/env python3
"""bench.py -- BN254 pairings/s on B200 (BASELINE.json metric), one JSON line on stdout (rank 0).

A "step" is one pass of the hot path over one batch of synthetic input: 2^16 independent BN254
pairings per GPU (BASELINE.json configs[1]; `--batch-log2` changes it).  Multi-GPU: one process per
GPU (torchrun), the batch index range is sharded with NO collective on the per-pairing path (weak
scaling: every rank gets its own 2^16 batch).

  value      pairings/s, inputs already resident in HBM, CUDA-event time on the launching stream
  e2e        same metric through the public host-buffer API (pinned host memory, H2D + D2H inside)
  roofline   dominant kernel (k_final_exp_hard_sync, its own CUDA-event time) against the integer-multiply
             (IMAD) pipe: algorithmic IMAD-equivalents (264 per Fq multiplication, lean
             Fq-mul counts) / event time, against the IMAD peak MEASURED in this run by a register-only
             microbenchmark (never below 64 IMAD/clk/SM at the max clock); executed Fq-mul counts beside it
  cpu_baseline  the C oracle (reference algorithm as written) on the box's host cores, bounded sample
  extra      BASELINE configs #3 and #4 (Groth16-shaped checks flat / with shared G2Precomp, G1 * Fr, G2 * Fr,
             their normalising window-kernel twins), each with device time, IMAD fraction, e2e and a CPU baseline;
             configs[0] latencies (n = 1 pairing, the Joux flow)
  sharded_product  BASELINE configs[4]'s exchange step: ONE product of 2^22 pairs split over the N ranks, 384-byte
             partials gathered, one final exponentiation; known answer one; pairs/s and the gather+combine latency
  single_call_multi_gpu  (N > 1) rank 0 drives all N GPUs through ONE bn254_pairing_many_sharded call

`--impl reference` times the reference's CPU algorithm (the C restatement: no Rust toolchain exists
in this image) on the same metric/config.
"""
import argparse
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
os.environ.setdefault("BN254_ORACLE_NATIVE", "1")   # the CPU legs time a -march=native build made on this machine

FQMUL_IMAD_EQ = 264                       # 128 full products x 2 + 8 low products
FQMUL_PRECOMP, FQMUL_MLOOP = 2916, 6045   # lean Fq-mul: AffineG2::precompute, G2Precomp::miller_loop
FQMUL_MILLER = FQMUL_PRECOMP + FQMUL_MLOOP
FQMUL_EASY, FQMUL_HARD = 228, 8541        # the reference's addition chain (plain square-and-multiply exp_by_neg_z)
FQMUL_FINALEXP = FQMUL_EASY + FQMUL_HARD
FQMUL_PAIRING = FQMUL_MILLER + FQMUL_FINALEXP   # 17,730
# what the kernels execute: width-4 signed-window exp_by_neg_z = 62 cyclotomic squarings (18) + 16 multiplications (54)
# three times, + 3 cyclotomic squarings + 10 multiplications + 3 Frobenius maps (21) in the chain (DESIGN F5)
FQMUL_HARD_EXECUTED = 3 * (62 * 18 + 16 * 54) + 3 * 18 + 10 * 54 + 3 * 21
FQMUL_GROTH16_FLAT = 4 * FQMUL_PRECOMP + 17268 + FQMUL_FINALEXP      # 37,701: all four line tables redone (as the reference does)
FQMUL_GROTH16_SHARED = FQMUL_PRECOMP + 17268 + FQMUL_FINALEXP        # 28,953: three shared G2Precomp tables
G1, G2, GT, FR = 96, 192, 384, 32
R_ORDER = 0x30644e72e131a029b85045b68181585d2833e84879b9709143e1f593f0000001
try:
    with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as _f:
        HBM_PEAK_GBS = json.load(_f).get("hbm_gbs")
except Exception:
    HBM_PEAK_GBS = 6538.6   # fallback: the figure the driver measured on this pool (B200_PROFILING.md)


def source_sha16():
    """Hash of the CUDA sources the library is built from: stamps profiles/ncu_traffic.json (tools/ncu_traffic.py writes it
    next to the capture) so that a traffic figure is only quoted for the build it was measured on."""
    h = hashlib.sha256()
    d = os.path.join(ROOT, "paritytech-bn_b200", "csrc")
    for fn in sorted(os.listdir(d)):
        if fn.endswith((".cu", ".cuh")):
            h.update(fn.encode())
            h.update(open(os.path.join(d, fn), "rb").read())
    return h.hexdigest()[:16]


def load_traffic():
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            t = json.load(f)
    except Exception:
        return None, "profiles/ncu_traffic.json missing"
    if t.get("source_sha16") != source_sha16():
        return None, "profiles/ncu_traffic.json was captured on another build (source_sha16 %s, this build %s): not quoted" % (
            t.get("source_sha16"), source_sha16())
    return t, "dram__bytes_read.sum + dram__bytes_write.sum per unit from the committed ncu --set full capture of this very build (%s)" % t.get("capture", "profiles/")


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch-log2", type=int, default=16, help="pairings per GPU per step = 2^k")
    ap.add_argument("--variant", default=None, help="library variant suffix (ad hoc A/B builds)")
    ap.add_argument("--cpu-sample", type=int, default=4096, help="pairings in the CPU-baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip configs #3/#4, the latency figures and the sharded product")
    ap.add_argument("--extra", action="store_true", help="(default) also time Groth16 checks, scalar muls, latencies, the sharded product")
    ap.add_argument("--sharded-log2", type=int, default=22, help="pairs of the ONE sharded product (BASELINE configs[4]: 2^22)")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1])); pw.append(float(r[2]))
                for nme, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(nme)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


def reference_arm(args, rank, world):
    """The reference's CPU implementation of the path (C restatement, all host threads), rank 0 only."""
    if rank != 0:
        return
    from oracle import c_oracle as c
    from tests import util
    cores = c.num_procs()
    # Bounded sample: the arm has to end within minutes on any host (the oracle does 0.35 k pairings/s per core, so the whole
    # 2^16 batch would take 12-25 s per step); throughput is flat in the batch size (independent units, static OpenMP split),
    # so 128 pairings per core per step measure the same rate.
    sample = min(1 << args.batch_log2, max(256, 128 * cores))
    g1, g2 = util.points_cpu(util.scalars(sample, stream=1), util.scalars(sample, stream=2))
    for _ in range(min(args.warmup, 1)):
        c.pairing_many(g1[:64 * G1], g2[:64 * G2], 0)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        c.pairing_many(g1, g2, 0)
    dt = time.perf_counter() - t0
    val = sample * args.steps / dt
    desc = ("%d pairings per step = 128 per host thread (bounded sample of the 2^%d batch: the CPU rate is flat in the batch size, "
            "the full batch would cost ~%.0f s per step), %d OpenMP threads, gcc -O3 -march=%s") % (
                sample, args.batch_log2, (1 << args.batch_log2) / val, cores, c.lib().build_flavour)
    print(json.dumps({
        "impl": "reference", "metric": "BN254 pairings/sec", "value": val, "unit": "pairings/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u32 limbs (256-bit Montgomery, mod q)", "data": "synthetic",
        "config": {"workload": "2^%d independent BN254 pairings per GPU (BASELINE configs[1]); CPU arm runs a bounded sample" % args.batch_log2},
        "cpu_baseline": {"value": val, "unit": "pairings/s", "cores": cores, "kind": "port", "sample": desc},
        "e2e": {"value": val, "unit": "pairings/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference crate itself cannot be built here (no cargo/rustc); this is the C restatement of its algorithm as written",
    }))


def hamming_work(fr_images, dbl, add):
    """SURVEY 8d count of the reference's double-and-add over canonical scalars: dbl (L - 1) + add (H - 1), summed exactly."""
    tot = 0
    rinv = pow(1 << 256, -1, R_ORDER)
    for i in range(0, len(fr_images), FR):
        k = int.from_bytes(fr_images[i:i + FR].tobytes(), "little") * rinv % R_ORDER
        if k:
            tot += dbl * (k.bit_length() - 1) + add * (bin(k).count("1") - 1)
    return tot


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        reference_arm(args, rank, world)
        return 0

    # Under torchrun every rank gets its own slice of the host cores BEFORE torch / CUDA start their threads, so that the ranks'
    # copy and launch threads do not migrate over one another (recorded in config.cpu_affinity; BN254_BENCH_PIN=0 disables).
    affinity = None
    if world > 1 and os.environ.get("BN254_BENCH_PIN", "1") != "0" and hasattr(os, "sched_setaffinity"):
        try:
            cores_all = sorted(os.sched_getaffinity(0))
            per = len(cores_all) // world
            if per >= 1:
                mine = cores_all[local * per:(local + 1) * per]
                os.sched_setaffinity(0, mine)
                affinity = "%d-%d of %d host cores per rank" % (mine[0], mine[-1], len(cores_all))
        except OSError:
            affinity = None

    import torch
    import torch.distributed as dist

    import paritytech_bn_b200 as bn
    from tests import util

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    host_group = None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
        host_group = dist.new_group(backend="gloo")     # host-side barrier (no spinning kernel on the waiting GPUs)
    eng = bn.Engine(local, variant=args.variant)
    stream = torch.cuda.ExternalStream(eng.stream, device=dev)
    n = 1 << args.batch_log2
    do_extra = not args.no_extra

    def dptr(t):
        return t.data_ptr()

    def gen_points(fr1, fr2):
        """(k1_i * G1, k2_i * G2) on the device from Fr images (numpy), by the CUDA scalar-multiplication kernels."""
        m = len(fr1) // FR
        d1 = torch.from_numpy(fr1).to(dev); d2 = torch.from_numpy(fr2).to(dev)
        gen1 = torch.from_numpy(np.tile(util.G1_GEN, m)).to(dev); gen2 = torch.from_numpy(np.tile(util.G2_GEN, m)).to(dev)
        o1 = torch.empty(m * G1, dtype=torch.uint8, device=dev); o2 = torch.empty(m * G2, dtype=torch.uint8, device=dev)
        torch.cuda.synchronize()
        eng.g1_mul_dev(dptr(gen1), dptr(d1), dptr(o1), m)
        eng.g2_mul_dev(dptr(gen2), dptr(d2), dptr(o2), m)
        eng.synchronize()
        return o1, o2

    # ---- synthetic input, generated on the device by the CUDA scalar-multiplication kernels -------
    ks1 = util.scalars(n, stream=1000 + rank)
    ks2 = util.scalars(n, stream=2000 + rank)
    d_g1, d_g2 = gen_points(util.fr_bytes(ks1), util.fr_bytes(ks2))
    d_out = torch.empty(n * GT, dtype=torch.uint8, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)     # > 126 MB L2

    def l2_flush():
        with torch.cuda.stream(stream):
            flush.zero_()

    def step():
        eng.pairing_dev(dptr(d_g1), dptr(d_g2), dptr(d_out), n)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def host_barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier(group=host_group)

    def timed_dev(fn, steps, warm=3):
        """Device time of `steps` calls of fn on the library's stream (L2 flushed before each), ms per call."""
        for _ in range(warm):
            l2_flush(); fn()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        torch.cuda.synchronize()
        for a, b in evs:
            l2_flush()
            a.record(stream); fn(); b.record(stream)
        torch.cuda.synchronize()
        return sum(a.elapsed_time(b) for a, b in evs) / steps

    def timed_host(fn, steps, warm=1):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(steps):
            fn()
        return (time.perf_counter() - t0) / steps

    def pinned(t):
        h = torch.empty(t.numel(), dtype=torch.uint8, pin_memory=True)
        h.copy_(t)
        return h

    for _ in range(max(args.warmup, 3)):
        l2_flush(); step()
    barrier()

    # ---- parity gate on a sample (never timed) ----------------------------------------------------------
    parity_n = 0
    if rank == 0:
        from oracle import c_oracle as c
        parity_n = 128
        h_g1 = d_g1[:parity_n * G1].cpu().numpy(); h_g2 = d_g2[:parity_n * G2].cpu().numpy()
        exp = c.pairing_many(h_g1, h_g2, 0)
        got = d_out[:parity_n * GT].cpu().numpy()
        if got.tobytes() != exp.tobytes():
            raise SystemExit("PARITY FAILURE: CUDA pairings differ from the oracle")

    # ---- timed region: K steps, device time on the launching stream -------------------------------------
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    launches0 = eng.launch_count()
    barrier()
    for k in range(args.steps):
        l2_flush()
        ev[k][0].record(stream)
        step()
        ev[k][1].record(stream)
    barrier()
    launches = eng.launch_count() - launches0
    total_ms = sum(a.elapsed_time(b) for a, b in ev)
    clocks = sampler.stop() if rank == 0 else None

    # ---- per-kernel split for the roofline: the profiling twin of the same entry point (identical launches,
    # each kernel bracketed by CUDA events on the library's stream) ------------------------------------------------
    best = None
    for _ in range(3):
        l2_flush()
        cur = eng.pairing_timed_dev(dptr(d_g1), dptr(d_g2), dptr(d_out), n)
        if best is None or sum(cur) < sum(best):
            best = cur
    t_aff, t_lines, t_mtab, t_easy, t_hard = best

    # ---- e2e: public host-buffer API, pinned host memory, H2D + D2H inside the timed region ----------------------
    h_g1, h_g2 = pinned(d_g1), pinned(d_g2)
    torch.cuda.synchronize()
    np_g1, np_g2 = h_g1.numpy(), h_g2.numpy()
    h_out = torch.empty(n * GT, dtype=torch.uint8, pin_memory=True)
    np_out = h_out.numpy()
    eng.pairing(np_g1, np_g2, out=np_out)
    barrier()
    e2e_steps = max(3, min(args.steps, 5))
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        res = eng.pairing(np_g1, np_g2, out=np_out)
    e2e_s = (time.perf_counter() - t0) / e2e_steps
    barrier()

    # ---- IMAD peak measured here (register-only microbenchmarks) --------------------------------------------------
    sms = torch.cuda.get_device_properties(local).multi_processor_count
    micro = {}
    for kind, nm in ((0, "imad"), (1, "imad_wide"), (2, "fq_mul")):
        ms, ops = eng.microbench(kind, sms * 8, 256, 2000 if kind < 2 else 400)
        micro[nm] = ops / (ms * 1e-3)

    # ---- aggregate over ranks ---------------------------------------------------------------------------------------
    t = torch.tensor([total_ms, e2e_s, t_aff, t_lines, t_mtab, t_easy, t_hard], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms, e2e_s, t_aff, t_lines, t_mtab, t_easy, t_hard = [float(x) for x in t.tolist()]
    t_miller = t_aff + t_lines + t_mtab

    value = n * world * args.steps / (total_ms * 1e-3)
    e2e_value = n * world / e2e_s
    # IMAD-equivalent peak: a wide multiply counts 2; use the better of the two measured forms
    peak_measured = max(micro["imad"], 2.0 * micro["imad_wide"])
    sm_mhz = (clocks or {}).get("sm_mhz") or 0.0
    # The microbenchmark runs right after the timed steps and can be power-capped on its own (seen at 2^18: 16.0 instead of
    # 18.5 T/s under sw_power_cap); a throttled denominator would flatter frac, so never go below the issue-rate figure
    # 64 IMAD/clk/SM at the maximum SM clock.
    sm_max_mhz = (clocks or {}).get("sm_max_mhz") or 0.0
    peak_live = peak_measured
    peak_measured = max(peak_measured, sms * 64 * sm_max_mhz * 1e6)
    if world > 1:   # every rank needs the same denominator for its extras
        pk = torch.tensor([peak_measured], dtype=torch.float64, device=dev)
        dist.broadcast(pk, src=0)
        peak_measured = float(pk.item())

    def frac_of(fqmul_total, ms):
        return fqmul_total * FQMUL_IMAD_EQ / (ms * 1e-3) / peak_measured

    out = None
    if rank == 0:
        # kernel names follow the library's path switches (defaults: slot-file kernels, one phase-synchronised block per SM)
        fexp_mode = os.environ.get("BN254_FEXP_SMEM", "2")
        hard_name = {"2": "k_final_exp_hard_sync", "1": "k_final_exp_hard_s"}.get(fexp_mode, "k_final_exp_many")
        mtab_name = "k_miller_table_affine" if os.environ.get("BN254_MILLER_SMEM", "1") == "0" else "k_miller_table_sync"
        lines_name = "k_g2_lines_affine" if os.environ.get("BN254_LINES_SMEM", "1") == "0" else "k_g2_lines_sync"
        # dominant kernel = the longest launch of the step, ONE kernel with its own event time
        cands = [(hard_name, t_hard, FQMUL_HARD if t_easy else FQMUL_FINALEXP, FQMUL_HARD_EXECUTED), (mtab_name, t_mtab, FQMUL_MLOOP, FQMUL_MLOOP),
                 (lines_name, t_lines, FQMUL_PRECOMP, FQMUL_PRECOMP)]
        if t_lines == 0.0:
            cands = cands[:1] + [("k_miller_many", t_mtab, FQMUL_MILLER, FQMUL_MILLER)]
        dom_name, dom_ms, dom_fq, dom_exec = max(cands, key=lambda cnd: cnd[1])
        achieved = dom_fq * FQMUL_IMAD_EQ * n / (dom_ms * 1e-3)
        table_bytes = 87 * 192 * n
        traffic_tab, traffic_note = load_traffic()
        traffic = (traffic_tab or {}).get("bytes_per_unit", {}).get(dom_name)
        algo_bytes = {hard_name: 2 * GT * n, mtab_name: table_bytes + (64 + GT) * n, lines_name: table_bytes + 128 * n}.get(dom_name)
        out = {
            "metric": "BN254 pairings/sec", "value": value, "unit": "pairings/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u32 limbs (256-bit Montgomery, mod q)", "data": "synthetic",
            "config": {"workload": "2^%d independent BN254 pairings per GPU (BASELINE configs[1]), random Jacobian points z != 1" % args.batch_log2,
                       "batch_per_gpu": n, "l2": "flushed between steps (256 MiB write)", "library": eng.build_info(),
                       "source_sha16": source_sha16(), "parity_checked_pairings": parity_n, "cpu_affinity": affinity},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "pairings/s", "h2d_bytes_per_step": n * (G1 + G2), "d2h_bytes_per_step": n * GT,
                    "api": "Engine.pairing -> bn254_pairing_many (pinned host buffers)", "ms_per_step": e2e_s * 1e3},
            "gpu_launches": int(launches),
            "roofline": {"bound": "imad", "kernel": dom_name, "kernel_ms": dom_ms, "achieved": achieved / 1e12, "peak": peak_measured / 1e12,
                         "unit": "TIMAD-eq/s", "frac": achieved / peak_measured,
                         "algorithmic_fq_mul": dom_fq, "executed_fq_mul": dom_exec,
                         "frac_executed": dom_exec * FQMUL_IMAD_EQ * n / (dom_ms * 1e-3) / peak_measured,
                         "count_note": "frac credits the reference's addition chain (SURVEY 8d, %d Fq-mul per unit); the kernel executes %d "
                                       "(width-4 signed-window exp_by_neg_z), so frac_executed is the pipe-side reading" % (dom_fq, dom_exec),
                         "traffic": traffic * n if traffic else None, "traffic_note": traffic_note,
                         "algorithmic_bytes": algo_bytes,
                         "peak_live_microbench": peak_live / 1e12,
                         "peak_source": "max(register-only mad.lo microbenchmark of this run, 64 IMAD/clk/SM x SMs x max SM clock); an IMAD.WIDE counts 2 IMAD-eq (measured 4.1 clk/warp-inst "
                                        "vs 2.0 for IMAD, profiles/r01_pipe_microbench2.txt); MEASURED_PEAKS.json has no integer peak",
                         "hbm": ({"achieved": traffic * n / (dom_ms * 1e-3) / 1e9, "peak": HBM_PEAK_GBS, "unit": "GB/s",
                                  "frac": traffic * n / (dom_ms * 1e-3) / 1e9 / HBM_PEAK_GBS,
                                  "note": "DRAM traffic of that kernel from the stamped capture: HBM is not the bound"}
                                 if traffic else None),
                         "peak_theoretical": sms * 64 * sm_mhz * 1e6 / 1e12 if sm_mhz else None,
                         "algorithmic_imad_eq_per_pairing": FQMUL_PAIRING * FQMUL_IMAD_EQ,
                         "whole_step_frac": FQMUL_PAIRING * FQMUL_IMAD_EQ * n * world * args.steps / (total_ms * 1e-3) / (peak_measured * world),
                         "kernels_ms": {"k_pairs_to_affine": t_aff, lines_name: t_lines, mtab_name: t_mtab, "k_final_exp_easy": t_easy,
                                        hard_name: t_hard, "miller_total": t_miller},
                         "kernels_frac": {lines_name: frac_of(FQMUL_PRECOMP * n, t_lines) if t_lines else None,
                                          mtab_name: frac_of(FQMUL_MLOOP * n, t_mtab) if t_mtab else None,
                                          hard_name: frac_of(FQMUL_HARD * n, t_hard) if t_hard else None},
                         "g2precomp_stream_hbm_gbs": {"write_" + lines_name: table_bytes / (t_lines * 1e-3) / 1e9 if t_lines else None,
                                                      "read_" + mtab_name: table_bytes / (t_mtab * 1e-3) / 1e9 if t_mtab else None,
                                                      "bytes": table_bytes, "hbm_peak_gbs": HBM_PEAK_GBS},
                         "microbench_ops_per_s": micro},
        }
        if not args.no_cpu_baseline:
            from oracle import c_oracle as c
            cores = c.num_procs()
            sample = min(n, args.cpu_sample)
            hs1, hs2 = np_g1[:sample * G1], np_g2[:sample * G2]
            c.pairing_many(hs1[:64 * G1], hs2[:64 * G2], 0)
            t0 = time.perf_counter()
            exp = c.pairing_many(hs1, hs2, 0)
            dt = time.perf_counter() - t0
            if exp.tobytes() != res[:sample * GT].tobytes():
                raise SystemExit("PARITY FAILURE: e2e CUDA pairings differ from the oracle on the CPU-baseline sample")
            if world == 1:      # the CPU baseline is a rank-0, N = 1 figure; at N > 1 the same sample only serves as the parity check
                out["cpu_baseline"] = {"value": sample / dt, "unit": "pairings/s", "cores": cores, "kind": "port",
                                       "sample": "first %d of the 2^%d pairings, C restatement of the reference algorithm as written (gcc -O3 -march=%s), OpenMP over all "
                                                 "host cores; outputs byte-compared with the CUDA results" % (sample, args.batch_log2, c.lib().build_flavour)}
            out["config"]["parity_checked_pairings"] = parity_n + sample

    # =====================================================================================================
    # extras: BASELINE configs #3, #4, #1 (rank 0) and the sharded product of configs[4] (all ranks)
    # =====================================================================================================
    if do_extra:
        extra = {}
        if rank == 0:
            try:
                extra.update(extras_rank0(args, eng, util, torch, dev, stream, timed_dev, timed_host, pinned, frac_of, dptr))
            except Exception as e:   # an extra must never cost the headline line
                extra["error"] = "%s: %s" % (type(e).__name__, e)
        del d_g1, d_g2, d_out, h_g1, h_g2, h_out
        torch.cuda.empty_cache()
        try:
            sp = sharded_product(args, eng, util, torch, dist, dev, rank, world, host_barrier, gen_points, dptr)
        except Exception as e:
            sp = {"error": "%s: %s" % (type(e).__name__, e)}
        if rank == 0:
            out["sharded_product"] = sp
            out["extra"] = extra
        if world > 1:
            try:
                sc = single_call_multi_gpu(args, bn, util, torch, dev, rank, world, host_barrier)
            except Exception as e:
                sc = {"error": "%s: %s" % (type(e).__name__, e)}
            if rank == 0:
                out["single_call_multi_gpu"] = sc
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        dist.barrier(group=host_group)
        dist.destroy_process_group()
    return 0


def extras_rank0(args, eng, util, torch, dev, stream, timed_dev, timed_host, pinned, frac_of, dptr):
    """BASELINE configs #3 (2^16 Groth16-shaped 4-pair checks, flat and with three shared G2Precomp tables), #4 (2^20 G1 * Fr,
    2^18 G2 * Fr: the reference-sequence kernels and the normalising window kernels) and #1 (n = 1 latency, Joux flow)."""
    from oracle import c_oracle as c
    cores = c.num_procs()
    ex = {}
    steps = max(3, min(args.steps, 5))

    # ---- config #3 -----------------------------------------------------------------------------------------
    nchk = 1 << 16

    def smul(which, ks):
        fr = util.fr_bytes(ks)
        if which == "g1":
            return eng.g1_mul(np.tile(util.G1_GEN, len(ks)), fr)
        return eng.g2_mul(np.tile(util.G2_GEN, len(ks)), fr)

    w = util.groth16_checks(nchk, stream=7, scalar_mul=smul)
    hp = {k: torch.from_numpy(w[k]).pin_memory() for k in ("g1", "g2", "g1_shared", "g1_var", "g2_var")}
    dv = {k: v.to(dev) for k, v in hp.items()}
    d_gt = torch.empty(nchk * GT, dtype=torch.uint8, device=dev)
    d_one = torch.empty(nchk, dtype=torch.uint8, device=dev)
    h_gt = torch.empty(nchk * GT, dtype=torch.uint8, pin_memory=True)
    h_one = torch.empty(nchk, dtype=torch.uint8, pin_memory=True)
    pc = eng.g2_precompute(w["vk_g2"])
    sample = 1024
    t0 = time.perf_counter()
    cg, cone = c.pairing_product_many(w["g1"][:sample * 4 * G1], w["g2"][:sample * 4 * G2], 4)
    cpu_rate = sample / (time.perf_counter() - t0)
    cpu3 = {"value": cpu_rate, "unit": "checks/s", "cores": cores, "kind": "port",
            "sample": "first %d of the 2^16 checks through the oracle's pairing_batch; Gt and accept bytes compared with the CUDA results" % sample}
    for name, fq, dev_fn, host_fn, h2d in (
            ("groth16_flat", FQMUL_GROTH16_FLAT,
             lambda: eng.pairing_batch_dev(dptr(dv["g1"]), dptr(dv["g2"]), 4, nchk, dptr(d_gt), dptr(d_one)),
             lambda: eng.pairing_batch(hp["g1"].numpy(), hp["g2"].numpy(), 4, out=h_gt.numpy(), is_one=h_one.numpy()),
             nchk * 4 * (G1 + G2)),
            ("groth16_shared_precomp", FQMUL_GROTH16_SHARED,
             lambda: eng.pairing_batch_precomp_dev(pc, dptr(dv["g1_shared"]), dptr(dv["g1_var"]), dptr(dv["g2_var"]), 1, nchk, dptr(d_gt), dptr(d_one)),
             lambda: eng.pairing_batch_precomp(pc, hp["g1_shared"].numpy(), hp["g1_var"].numpy(), hp["g2_var"].numpy(), 1,
                                               out=h_gt.numpy(), is_one=h_one.numpy()),
             nchk * (4 * G1 + G2))):
        ms = timed_dev(dev_fn, steps)
        got_gt, got_one = d_gt[:sample * GT].cpu().numpy(), d_one.cpu().numpy()
        if got_gt.tobytes() != cg.tobytes() or (got_one != w["valid"]).any():
            raise SystemExit("PARITY FAILURE: %s differs from the oracle" % name)
        e2e_s = timed_host(host_fn, steps)
        if (h_one.numpy() != w["valid"]).any():
            raise SystemExit("PARITY FAILURE: %s (host path) accept bytes" % name)
        ex[name] = {"value": nchk / (ms * 1e-3), "unit": "checks/s", "ms": ms, "n_checks": nchk, "pairs_per_check": 4,
                    "algorithmic_fq_mul_per_check": fq, "algorithmic_imad_eq": fq * FQMUL_IMAD_EQ * nchk, "frac": frac_of(fq * nchk, ms),
                    "e2e": {"value": nchk / e2e_s, "unit": "checks/s", "h2d_bytes": h2d, "d2h_bytes": nchk * (GT + 1)},
                    "cpu_baseline": cpu3, "accept_reject_ok": True}
    pc.close()
    del dv, hp, d_gt, d_one, h_gt, h_one

    # ---- config #4 -----------------------------------------------------------------------------------------
    for which, log2n, unit, dbl, add, sample in (("g1", 20, G1, 7, 16, 1 << 14), ("g2", 18, G2, 16, 43, 1 << 12)):
        m = 1 << log2n
        gen = util.G1_GEN if which == "g1" else util.G2_GEN
        mul_dev, muln_dev = (eng.g1_mul_dev, eng.g1_mul_normalized_dev) if which == "g1" else (eng.g2_mul_dev, eng.g2_mul_normalized_dev)
        mul_h, muln_h = (eng.g1_mul, eng.g1_mul_normalized) if which == "g1" else (eng.g2_mul, eng.g2_mul_normalized)
        cmul, cnorm = (c.g1_mul, c.g1_normalize) if which == "g1" else (c.g2_mul, c.g2_normalize)
        fr0, fr = util.fr_images(m, stream=301), util.fr_images(m, stream=302)
        fr[:4 * FR] = util.fr_bytes([0, 1, 2, R_ORDER - 1])
        d_gen = torch.from_numpy(np.tile(gen, m)).to(dev)
        d_fr0, d_fr = torch.from_numpy(fr0).to(dev), torch.from_numpy(fr).to(dev)
        d_base = torch.empty(m * unit, dtype=torch.uint8, device=dev)
        d_res = torch.empty(m * unit, dtype=torch.uint8, device=dev)
        torch.cuda.synchronize()
        mul_dev(dptr(d_gen), dptr(d_fr0), dptr(d_base), m)          # bases k_i * G, Jacobian with z != 1
        eng.synchronize()
        del d_gen
        h_base, h_fr = pinned(d_base), torch.from_numpy(fr).pin_memory()
        h_res = torch.empty(m * unit, dtype=torch.uint8, pin_memory=True)
        # the section-8d count is exact over the whole batch only through Python big integers: take it on a 2^14 sample
        cnt_n = 1 << 14
        work = hamming_work(fr[:cnt_n * FR], dbl, add) / cnt_n
        t0 = time.perf_counter()
        craw = cmul(h_base.numpy()[:sample * unit], fr[:sample * FR])
        cpu_rate = sample / (time.perf_counter() - t0)
        cnrm = cnorm(craw)
        cpu4 = {"value": cpu_rate, "unit": "scalar muls/s", "cores": cores, "kind": "port",
                "sample": "first %d of the 2^%d (point, scalar) pairs; raw Jacobian and normalised bytes compared with the CUDA results" % (sample, log2n)}
        for name, dfn, hfn, expect in (
                ("%s_mul" % which, lambda: mul_dev(dptr(d_base), dptr(d_fr), dptr(d_res), m),
                 lambda: mul_h(h_base.numpy(), h_fr.numpy(), out=h_res.numpy()), craw),
                ("%s_mul_normalized" % which, lambda: muln_dev(dptr(d_base), dptr(d_fr), dptr(d_res), m),
                 lambda: muln_h(h_base.numpy(), h_fr.numpy(), out=h_res.numpy()), cnrm)):
            ms = timed_dev(dfn, steps)
            if d_res[:sample * unit].cpu().numpy().tobytes() != expect.tobytes():
                raise SystemExit("PARITY FAILURE: %s differs from the oracle" % name)
            e2e_s = timed_host(hfn, steps)
            ex[name] = {"value": m / (ms * 1e-3), "unit": "scalar muls/s", "ms": ms, "n": m,
                        "algorithmic_fq_mul_per_unit": work, "algorithmic_imad_eq": work * FQMUL_IMAD_EQ * m, "frac": frac_of(work * m, ms),
                        "count": "%d (L - 1) + %d (H - 1) Fq-mul over the canonical scalars (SURVEY 8d), mean of a 2^14 sample" % (dbl, add),
                        "e2e": {"value": m / e2e_s, "unit": "scalar muls/s", "h2d_bytes": m * (unit + FR), "d2h_bytes": m * unit},
                        "cpu_baseline": cpu4}
            if name.endswith("_normalized"):
                ex[name]["frac_note"] = ("credited with the reference's double-and-add count above; the signed 4-bit window executes "
                                         "fewer multiplications (DESIGN F16), so the credited fraction can exceed 1 -- it is not a pipe utilisation")
        del d_base, d_res, d_fr0, d_fr, h_base, h_fr, h_res

    # ---- config #1: latency of one pairing and of the Joux flow (examples/joux.rs:7-21) ------------------------
    a, b, cc = util.scalars(3, stream=21)
    fa, fb, fc = (util.fr_bytes([x]) for x in (a, b, cc))
    pa, pb, pc_ = (eng.g1_mul(util.G1_GEN, f) for f in (fa, fb, fc))
    qa, qb, qc = (eng.g2_mul(util.G2_GEN, f) for f in (fa, fb, fc))
    lat1 = min(timed_host(lambda: eng.pairing(pa, qb), 5, warm=2) for _ in range(3))
    g1j, g2j, frj = np.concatenate([pb, pc_, pa]), np.concatenate([qc, qa, qb]), np.concatenate([fa, fb, fc])

    def joux():
        e = eng.pairing(g1j, g2j)
        return eng.gt_pow(e, frj)
    latj = min(timed_host(joux, 5, warm=2) for _ in range(3))
    k = joux().reshape(3, GT)
    if not (k[0].tobytes() == k[1].tobytes() == k[2].tobytes()) or eng.pairing(pa, qb).tobytes() != c.pairing_many(pa, qb, 1).tobytes():
        raise SystemExit("PARITY FAILURE: Joux keys differ / n = 1 pairing differs from the oracle")
    t0 = time.perf_counter()
    for _ in range(4):
        c.pairing_many(pa, qb, 1)
    cpu1 = (time.perf_counter() - t0) / 4
    t0 = time.perf_counter()
    e = c.pairing_many(g1j, g2j, 1); c.gt_pow(e, frj)
    cpuj = time.perf_counter() - t0
    ex["latency"] = {"n1_pairing_ms": lat1 * 1e3, "joux_3_pairings_3_pow_ms": latj * 1e3,
                     "cpu_port_single_thread": {"n1_pairing_ms": cpu1 * 1e3, "joux_ms": cpuj * 1e3},
                     "api": "Engine.pairing / Engine.gt_pow with host buffers, wall clock around the call (copies and launch latency inside)",
                     "joux_keys_equal": True}
    return ex


def sharded_product(args, eng, util, torch, dist, dev, rank, world, host_barrier, gen_points, dptr):
    """BASELINE configs[4]: ONE product of 2^k pairs split over the ranks by contiguous index range.  Pair i < n/2 is
    (A_i G1, B_i G2); pair n/2 + j is ((-B_j) G1, A_j G2): the exponents cancel, so the whole product is exactly one while every
    rank's partial is an arbitrary Fq12 (the partner of a pair lives half the index range away, i.e. on another rank for even N).
    Each rank: Miller values of its range on the split path, tree product -> 384 bytes; the N partials are gathered (NCCL
    all_gather of 384 B over NVLink; nothing on the per-pairing path), multiplied in rank order and exponentiated ONCE."""
    n = 1 << args.sharded_log2
    half = n // 2
    A, B = util.fr_images(half, stream=401), util.fr_images(half, stream=402)
    lo, hi = n * rank // world, n * (rank + 1) // world

    f1, f2 = util.cancelling_pair_scalars(A, B, lo, hi)
    cnt = hi - lo
    d_g1, d_g2 = gen_points(f1, f2)
    d_part = torch.empty(GT, dtype=torch.uint8, device=dev)
    gathered = torch.empty(world * GT, dtype=torch.uint8, device=dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    stream = torch.cuda.ExternalStream(eng.stream, device=dev)
    result = {}
    for it in range(2):                       # first pass warms the arenas up
        host_barrier()
        t_start = time.perf_counter()
        e0.record(stream)
        eng.miller_product_dev(dptr(d_g1), dptr(d_g2), cnt, dptr(d_part))
        e1.record(stream)
        eng.synchronize()
        t_part_host = time.perf_counter()
        if world > 1:
            dist.all_gather_into_tensor(gathered, d_part)
        else:
            gathered.copy_(d_part)
        parts = gathered.cpu().numpy().reshape(world, GT)
        acc = parts[0].copy()
        for d in range(1, world):
            acc = eng.gt_mul(acc, parts[d])
        gt, some = eng.final_exponentiation(acc)
        t_end = time.perf_counter()
        ms_part = e0.elapsed_time(e1)
        tt = torch.tensor([ms_part, (t_end - t_part_host) * 1e3, (t_end - t_start) * 1e3], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms_part, ms_comb, ms_all = [float(x) for x in tt.tolist()]
        is_one = bool(some[0]) and gt.tobytes() == util.GT_ONE.tobytes()
        partial_is_one = parts[0].tobytes() == util.GT_ONE.tobytes()
        result = {"pairs": n, "ranks": world, "value": n / ((ms_part + ms_comb) * 1e-3), "unit": "pairs/s",
                  "partials_ms_max_over_ranks": ms_part, "gather_combine_us": ms_comb * 1e3, "wall_ms": ms_all,
                  "is_one": is_one, "rank0_partial_is_one": partial_is_one,
                  "gather": "all_gather of one 384-byte Fq12 per rank (NCCL over NVLink)" if world > 1 else "single rank: no gather",
                  "construction": "pair i: (A_i G1, B_i G2); pair n/2 + i: ((-B_i) G1, A_i G2) -- product exactly one, partials arbitrary",
                  "per_gpu_pairs_per_s": cnt / (ms_part * 1e-3)}
    if not result["is_one"]:
        raise SystemExit("PARITY FAILURE: the sharded product of cancelling pairs is not one")
    # tampered twin: one pair replaced on rank 0 must flip the answer (not timed)
    if rank == 0:
        d_g1[:G1] = torch.from_numpy(util.G1_GEN).to(dev)
    eng.miller_product_dev(dptr(d_g1), dptr(d_g2), cnt, dptr(d_part))
    eng.synchronize()
    if world > 1:
        dist.all_gather_into_tensor(gathered, d_part)
    else:
        gathered.copy_(d_part)
    parts = gathered.cpu().numpy().reshape(world, GT)
    acc = parts[0].copy()
    for d in range(1, world):
        acc = eng.gt_mul(acc, parts[d])
    gt, _ = eng.final_exponentiation(acc)
    result["tampered_is_one"] = gt.tobytes() == util.GT_ONE.tobytes()
    if result["tampered_is_one"]:
        raise SystemExit("PARITY FAILURE: a tampered sharded product still reads one")
    return result


def single_call_multi_gpu(args, bn, util, torch, dev, rank, world, host_barrier):
    """ONE bn254_pairing_many_sharded call from rank 0 over all N GPUs of the box (the other ranks wait on a host barrier, their
    GPUs idle): n independent pairs in one pinned host buffer, contiguous index ranges, one host thread per device."""
    host_barrier()
    res = None
    if rank == 0 and torch.cuda.device_count() >= world:
        k = min(args.sharded_log2, 16 + max(0, world.bit_length() - 1) + 1)      # 2^17 per GPU
        n = 1 << k
        base = 1 << 14
        engines = [bn.Engine(d) for d in range(world)]
        try:
            p1 = engines[0].g1_mul(np.tile(util.G1_GEN, base), util.fr_images(base, stream=501)).reshape(base, G1)
            p2 = engines[0].g2_mul(np.tile(util.G2_GEN, base), util.fr_images(base, stream=502)).reshape(base, G2)
            idx = (np.arange(n, dtype=np.int64) * 2654435761 % base)
            h1 = torch.from_numpy(p1[idx].reshape(-1)).pin_memory()
            h2 = torch.from_numpy(p2[(idx * 7 + 3) % base].reshape(-1)).pin_memory()
            ho = torch.empty(n * GT, dtype=torch.uint8, pin_memory=True)
            bn.pairing_sharded(engines, h1.numpy(), h2.numpy(), out=ho.numpy())          # warm-up (arenas)
            ts = []
            for _ in range(5):
                t0 = time.perf_counter()
                bn.pairing_sharded(engines, h1.numpy(), h2.numpy(), out=ho.numpy())
                ts.append(time.perf_counter() - t0)
            dt = sorted(ts)[len(ts) // 2]                                                  # median of five calls
            from oracle import c_oracle as c
            sel = np.concatenate([np.arange(8), np.arange(n // 2 - 4, n // 2 + 4), np.arange(n - 8, n)])
            exp = c.pairing_many(h1.numpy().reshape(n, G1)[sel].reshape(-1), h2.numpy().reshape(n, G2)[sel].reshape(-1), 0)
            ok = ho.numpy().reshape(n, GT)[sel].tobytes() == exp.tobytes()
            if not ok:
                raise SystemExit("PARITY FAILURE: bn254_pairing_many_sharded differs from the oracle")
            res = {"pairings": n, "devices": world, "value": n / dt, "unit": "pairings/s", "ms": dt * 1e3,
                   "ms_min": min(ts) * 1e3, "calls": len(ts),
                   "api": "bn254_pairing_many_sharded(ctxs[%d], ...) from ONE process, pinned host buffers, copies inside; median of the calls" % world,
                   "parity_sample_ok": ok}
        finally:
            for e in engines:
                e.close()
    host_barrier()
    return res


if __name__ == "__main__":
    sys.exit(main())
