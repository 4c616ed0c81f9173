#!/usr/bin/env python
"""bench.py -- minnow encode/decode hot path on B200 (driver contract).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

One "step" = one encode pass + one decode pass of the hot path over one
snapshot slab of synthetic Zel'dovich-displaced Lagrangian-grid particles:
per GPU 1024^3 particles, positions + velocities, planar float32 (24 GiB raw),
cut into 256 segments of 2^22 consecutive particle ids = 512 independent
fields. At N=1 that is BASELINE.json configs[2]; at
N>1 rank r holds ids [r*2^30, (r+1)*2^30) of a 2048^3 box (configs[3] at
N=8), weak scaling, no data-path collective: the only exchange is the
all-gather of per-rank compressed sizes for the output offset prefix, which is
inside the timed region.

metric/value: GB/s of raw float32 input through the codec, encode and decode
passes together: value = N * 2 * raw_bytes / (t_encode + t_decode), inputs
resident in HBM. `e2e` is the same metric through the host-buffer C-ABI calls
(mnwcuda_EncodeFieldsHost / mnwcuda_DecodeFieldsHost) from pinned host memory.

--impl reference times the reference's own CPU primitives (oracle/_ref, the
reference's util.c/rand.c/quant.c compiled from /root/reference by
oracle/Makefile; else the restated oracle port) composed into the same
algorithm, on all host cores, on a bounded sample of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SEG_LOG2 = 22
SEG = 1 << SEG_LOG2
BOX = 100.0
POS_DELTA = 1e-3   # getSegment(), /root/reference/src/funcs.c:235
VEL_DELTA = 1.0    # /root/reference/src/funcs.c:247
GEN_SEED = 20260401
METRIC = "encode+decode GB/s raw input"
UNIT = "GB/s"


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# ---------------------------------------------------------------------------
# clocks sampler (B200_PROFILING.md: clocks DURING the timed region)
# ---------------------------------------------------------------------------
class Clocks:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.lines:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
                pw.append(float(parts[2]))
            except ValueError:
                continue
            for name, val in zip(names, parts[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "power_w_max": float(max(pw)),
                "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------
# workload
# ---------------------------------------------------------------------------
FIELD_POSN, FIELD_VELC = 0x506F736E, 0x56656C63


def field_seed(i, first_field=0):
    """Dither seed of field i of the snapshot: every field its own (the reference
    seeds every undoFloat call afresh, /root/reference/src/quant.c:639)."""
    return 0x5EED0000 + first_field + i


def field_list(n_segments, seg=SEG, first_field=0, same_seed=False):
    """[(float offset, n, code, delta, width, seed)]: field 2s = positions of
    segment s, 2s+1 = velocities; each 3*seg planar floats."""
    out = []
    for s in range(n_segments):
        sp = 0 if same_seed else field_seed(2 * s, first_field)
        sv = 0 if same_seed else field_seed(2 * s + 1, first_field)
        out.append(((2 * s) * 3 * seg, seg, FIELD_POSN, POS_DELTA, BOX, sp))
        out.append(((2 * s + 1) * 3 * seg, seg, FIELD_VELC, VEL_DELTA, 0.0, sv))
    return out


def generate(codec, d_in, ngrid, first_id, n_segments, seg=SEG, noise_cells=0.0):
    for s in range(n_segments):
        pos = d_in[(2 * s) * 3 * seg:(2 * s + 1) * 3 * seg]
        vel = d_in[(2 * s + 1) * 3 * seg:(2 * s + 2) * 3 * seg]
        codec.generate(ngrid, BOX, GEN_SEED, first_id + s * seg, seg, pos, vel, noise_cells=noise_cells)


P2P_COPY_CEILING_GBS = 774.5  # scripts/p2p_peak.py on a 4 x B200 box of this pool
NOMINAL_HBM_GBS = 8000.0  # HBM3e as specified; the roofline's denominator is the measured copy bandwidth


def native_so_loaded():
    """In-tree shared objects mapped into this process (from /proc/self/maps)."""
    seen = []
    try:
        with open("/proc/self/maps") as fh:
            for ln in fh:
                path = ln.rstrip("\n").split(None, 5)[-1] if len(ln.split(None, 5)) == 6 else ""
                if path.endswith(".so") and os.path.realpath(path).startswith(os.path.realpath(ROOT) + os.sep):
                    rel = os.path.relpath(os.path.realpath(path), os.path.realpath(ROOT))
                    if rel not in seen:
                        seen.append(rel)
    except OSError:
        pass
    return sorted(seen)


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ---------------------------------------------------------------------------
# CPU arm: the reference's primitives on host cores
# ---------------------------------------------------------------------------
def load_oracle():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import _oracle
    ref = _oracle.reference()
    if ref is not None:
        return ref, "reference"
    return _oracle.port(), "port"


def host_sample(orc, n_seg, seg, ngrid, threads):
    """n_seg segments of pos+vel made on the host by the oracle's generator (the
    device generator's recipe on host libm): the reference arm never loads the
    CUDA library. Returns (pos[n_seg*3*seg], vel[...])."""
    pos = np.empty(n_seg * 3 * seg, np.float32)
    vel = np.empty(n_seg * 3 * seg, np.float32)
    for s in range(n_seg):
        p, v = orc.gen_zeldovich(ngrid, BOX, GEN_SEED, s * seg, seg, 0.0, threads)
        pos[s * 3 * seg:(s + 1) * 3 * seg] = p
        vel[s * 3 * seg:(s + 1) * 3 * seg] = v
    return pos, vel


def cpu_baseline(orc, pos, vel, threads, n_seg, seg, seeds=None):
    """Stepper over a host sample: one step = encode + decode of n_seg pos fields
    and n_seg vel fields, independent fields spread over `threads` host threads."""
    payloads = []
    if seeds is None:
        seeds = [field_seed(i) for i in range(2 * n_seg)]
    for s in range(n_seg):
        payloads.append(orc.encode_field(FIELD_POSN, pos[s * 3 * seg:(s + 1) * 3 * seg], POS_DELTA, BOX,
                                         seed=seeds[2 * s]))
        payloads.append(orc.encode_field(FIELD_VELC, vel[s * 3 * seg:(s + 1) * 3 * seg], VEL_DELTA, 0.0,
                                         seed=seeds[2 * s + 1]))
    raw = 2 * n_seg * 3 * seg * 4

    def one_step(o=orc, nthreads=threads, k=n_seg):
        te = 0.0
        for code, data, delta, width in ((FIELD_POSN, pos, POS_DELTA, BOX), (FIELD_VELC, vel, VEL_DELTA, 0.0)):
            secs, _ = o.bench_encode(code, data[:k * 3 * seg], seg, k, delta, width, nthreads)
            te += secs
        td = o.bench_decode(payloads[:2 * k], seg, nthreads)
        return te, td

    return raw, payloads, one_step


def extra_cpu_figures(orc, one_step, seg):
    """The two other figures BASELINE.md section 4 promises: 1 thread (what the
    reference can do by itself: it has no threads) and the -D DEBUG build its own
    Makefile ships (Makefile:13-15), on a 1-segment sample."""
    out = {}
    raw1 = 2 * 3 * seg * 4
    te, td = one_step(orc, 1, 1)
    out["threads1_gbs"] = 2 * raw1 / (te + td) / 1e9
    out["threads1_encode_gbs"] = raw1 / te / 1e9
    out["threads1_decode_gbs"] = raw1 / td / 1e9
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import _oracle
    dbg = _oracle.reference_debug()
    if dbg is not None:
        te, td = one_step(dbg, 1, 1)
        out["debug_gbs"] = 2 * raw1 / (te + td) / 1e9
        out["debug_note"] = "reference compiled -D DEBUG (its Makefile's flags), 1 thread, 1 segment"
    else:
        out["debug_gbs"] = None
    return out


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    threads = os.cpu_count() or 1
    n_seg = args.cpu_segments or max(1, min(threads, 32))
    threads = min(threads, n_seg)
    ngrid = 1024 if args.gpus == 1 else 2048
    orc, kind = load_oracle()
    pos, vel = host_sample(orc, n_seg, SEG, ngrid, os.cpu_count() or 1)
    raw, _payloads, one_step = cpu_baseline(orc, pos, vel, threads, n_seg, SEG)
    for _ in range(args.warmup):
        one_step()
    te = td = 0.0
    t0 = time.perf_counter()
    for _ in range(args.steps):
        a, b = one_step()
        te += a
        td += b
    wall = time.perf_counter() - t0
    value = 2 * raw * args.steps / (te + td) / 1e9
    sample = ("%d segments x 2^%d particles, pos+vel (%.0f MiB raw) per step, host generator (oracle); "
              "encode %.3f GB/s, decode %.3f GB/s"
              % (n_seg, SEG_LOG2, raw / 2**20, raw * args.steps / te / 1e9, raw * args.steps / td / 1e9))
    base = {"value": value, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample}
    base.update(extra_cpu_figures(orc, one_step, SEG))
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * (te + td) / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32/u32",
        "data": "synthetic",
        "config": workload_config(args.gpus, sample_only=True),
        "cpu_baseline": base,
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "wall_s": wall,
        # what this process mapped from the repository: the oracle's libraries, nothing of the product
        "native_so_loaded": native_so_loaded(),
    }
    print(json.dumps(line), flush=True)
    return 0


def workload_config(n_gpus, sample_only=False, **extra):
    ngrid = 1024 if n_gpus == 1 else 2048
    cfg = {
        "workload": ("1024^3 pos+vel Zel'dovich snapshot (BASELINE configs[2], 24 GiB raw)" if n_gpus == 1 else
                     "2048^3 box pos+vel, ids [r*2^30,(r+1)*2^30) per rank (BASELINE configs[3] slabs, 24 GiB raw per GPU)"),
        "ngrid": ngrid, "particles_per_gpu": 1 << 30, "segment_particles": SEG, "fields_per_gpu": 512,
        "box": BOX, "pos_delta": POS_DELTA, "vel_delta": VEL_DELTA, "gen_seed": GEN_SEED,
        "gen": "64 sine modes |k_i|<=8, rms displacement 1.5 cells, v=200*psi/cell",
        "l2": "inputs larger than L2 (24 GiB streamed per pass)",
    }
    cfg.update(extra)
    return cfg


# ---------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------
def run_b200(args):
    import torch
    import torch.distributed as dist

    import minnow_c_b200 as mc
    from minnow_c_b200 import sharded

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: libminnow_cuda has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    n_gpus = world

    n_seg = args.segments
    seg = SEG
    ngrid = 1024 if n_gpus == 1 else 2048
    first_id = rank * n_seg * seg
    codec = mc.Codec(local)
    fields = field_list(n_seg, seg, first_field=rank * 2 * n_seg, same_seed=args.same_seed)
    nf = len(fields)
    descs = mc.make_descs(fields)
    raw_bytes = nf * 3 * seg * 4

    d_in = torch.empty(nf * 3 * seg, dtype=torch.float32, device=dev)
    generate(codec, d_in, ngrid, first_id, n_seg, seg)
    cap = nf * mc.field_bound(seg)
    d_out = torch.empty(cap, dtype=torch.uint8, device=dev)
    d_off = torch.zeros(nf + 1, dtype=torch.int64, device=dev)
    d_dec = torch.empty_like(d_in)
    torch.cuda.synchronize()

    # N > 1: the C multi-GPU driver (libminnow_nccl.so, include/minnow_nccl.h); only
    # the ncclUniqueId travels through torch.distributed
    shard = None
    if world > 1:
        uid = [mc.Sharded.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        shard = mc.Sharded(codec, rank, world, uid[0])
        d_sizes = torch.zeros(world, dtype=torch.int64, device=dev)
        d_base = torch.zeros(1, dtype=torch.int64, device=dev)

    def step(record=None):
        if record:
            record[0].record()
        if shard is not None:
            # encode + output placement of the sharded stream in one C call: ncclAllGather
            # of the per-rank packed sizes on the encode stream, exclusive prefix on the device
            shard.encode(d_in, descs, nf, d_out, d_off, d_sizes, d_base)
        else:
            codec.encode_dev(d_in, descs, nf, d_out, d_off)
        if record:
            record[1].record()
        if args.encode_only:
            if record:
                record[2].record()
            return
        codec.decode_dev(d_out, d_off, descs, nf, d_dec)
        if record:
            record[2].record()

    for _ in range(max(args.warmup, 1)):
        step()
    torch.cuda.synchronize()
    if codec.status() != 0 and not args.encode_only:
        raise SystemExit("device-side error during warm-up")
    comp_bytes = int(d_off[nf].item())

    # ---- timed region: K steps, CUDA events on the launching stream ----
    codec.enable_kernel_timing(True)
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(args.steps)]
    ktimes = []
    clocks = Clocks(local)
    if rank == 0:
        clocks.start()
    launches0 = codec.launches
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t_wall = time.perf_counter()
    for i in range(args.steps):
        step(evs[i])
        if i == args.steps - 1:
            torch.cuda.synchronize()
            ktimes.append(codec.kernel_times())
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t_wall = time.perf_counter() - t_wall
    clk = clocks.stop() if rank == 0 else None
    launches = codec.launches - launches0
    codec.enable_kernel_timing(False)
    if codec.status() != 0 and not args.encode_only:
        raise SystemExit("device-side error during the timed region")
    if args.encode_only:
        if rank == 0:
            em = sum(e[0].elapsed_time(e[1]) for e in evs) / args.steps
            print(json.dumps({"encode_only_ms": em, "debug_mask": os.environ.get("MNWCUDA_DEBUG_MASK")}), flush=True)
        codec.close()
        return 0

    enc_ms = sum(e[0].elapsed_time(e[1]) for e in evs)
    dec_ms = sum(e[1].elapsed_time(e[2]) for e in evs)
    tot_ms = evs[0][0].elapsed_time(evs[-1][2])
    t = torch.tensor([enc_ms, dec_ms, tot_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    enc_ms, dec_ms, tot_ms = [float(v) for v in t.cpu()]

    # ---- size-independent property check (untimed): round-trip error bound ----
    check = roundtrip_check(torch, d_in, d_dec, fields, seg)

    # ---- e2e through the host-buffer C ABI (pinned host memory) ----
    e2e = None if args.no_e2e else run_e2e(torch, mc, codec, d_in, args, n_gpus, dev, world)

    gather = None
    if world > 1 and not args.no_gather:
        gather = run_gather(torch, dist, shard, d_out, d_sizes, d_base, dev, rank)
    shard_parity = None
    if world > 1 and not args.no_cpu:
        shard_parity = sharded_parity(torch, dist, d_in, d_out, d_off, fields, dev, pick=(0, 1, nf - 2, nf - 1))

    value = n_gpus * 2 * raw_bytes * args.steps / (tot_ms * 1e-3) / 1e9
    enc_gbs = n_gpus * raw_bytes * args.steps / (enc_ms * 1e-3) / 1e9
    dec_gbs = n_gpus * raw_bytes * args.steps / (dec_ms * 1e-3) / 1e9

    if rank != 0:
        if shard is not None:
            shard.close()
        if world > 1:
            dist.destroy_process_group()
        return 0

    # ---- roofline of the dominant kernel (live CUDA-event durations) ----
    peak, peak_src = peaks()
    kt = ktimes[-1]
    c = comp_bytes / (nf * 3 * seg)  # compressed bytes per value (payload + tables + headers)
    alg_bytes = (4.0 + c) * nf * 3 * seg  # per launch: SURVEY 8(d): encode 4 B read + c B written; decode mirror
    dom = "k_encode" if kt["k_encode"] >= kt["k_unpack_dequant"] else "k_unpack_dequant"
    roof = {}
    for name in ("k_encode", "k_unpack_dequant"):
        roof[name] = {"ms": kt[name], "achieved": alg_bytes / (kt[name] * 1e-3) / 1e9, "alg_bytes": alg_bytes}
    roofline = {
        "bound": "hbm", "kernel": dom, "achieved": roof[dom]["achieved"], "peak": peak, "unit": "GB/s",
        "frac": roof[dom]["achieved"] / peak, "traffic": load_traffic(dom), "peak_source": peak_src,
        # the same against the nominal HBM3e figure (every frac in this line is against `peak`;
        # multiply by peak / nominal_peak for the nominal one)
        "nominal_peak": NOMINAL_HBM_GBS, "frac_of_nominal": roof[dom]["achieved"] / NOMINAL_HBM_GBS,
        "alg_bytes_per_value": 4.0 + c, "values_per_launch": nf * 3 * seg,
        "kernels": {k: {"ms": round(v["ms"], 4), "achieved_gbs": round(v["achieved"], 1),
                        "frac": round(v["achieved"] / peak, 4),
                        "frac_of_nominal": round(v["achieved"] / NOMINAL_HBM_GBS, 4)} for k, v in roof.items()},
        "encode_pass": {"ms": enc_ms / args.steps,
                        "achieved_gbs": alg_bytes / (enc_ms / args.steps * 1e-3) / 1e9,
                        "frac": alg_bytes / (enc_ms / args.steps * 1e-3) / 1e9 / peak},
        "decode_pass": {"ms": dec_ms / args.steps,
                        "achieved_gbs": alg_bytes / (dec_ms / args.steps * 1e-3) / 1e9,
                        "frac": alg_bytes / (dec_ms / args.steps * 1e-3) / 1e9 / peak},
    }

    cpu = None
    if n_gpus == 1 and not args.no_cpu:
        offs_h = d_off.cpu().numpy()

        def gpu_payload(i):
            return d_out[int(offs_h[i]):int(offs_h[i + 1])].cpu().numpy()

        cpu = run_cpu_leg(args, d_in, fields, gpu_payload, nf)
        if cpu.get("parity") is not None:
            check["sample_bytes_vs_reference"] = cpu["parity"]
            check["ok"] = bool(check["ok"] and cpu["parity"]["ok"])
    if shard_parity is not None:
        check["sample_bytes_vs_reference"] = shard_parity
        check["ok"] = bool(check["ok"] and shard_parity["ok"])

    checksums = refflow = halves = None
    if n_gpus == 1 and not args.no_sweep:
        halves = run_encode_halves(torch, codec, d_in, descs, nf, d_out, d_off, alg_bytes, peak)
        checksums = run_checksums(torch, mc, codec, d_out, d_off, descs, nf, comp_bytes, peak)
        refflow = run_reference_flow(torch, mc, d_in, seg, fields)
    sweep = None
    if n_gpus == 1 and not args.no_sweep:
        sweep = run_sweep(torch, mc, codec, d_in, d_out, d_off, d_dec, fields, seg, peak, ngrid, first_id, n_seg)

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": n_gpus, "steps": args.steps,
        "warmup": max(args.warmup, 1), "ms_per_step": tot_ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32/u32", "data": "synthetic",
        "config": workload_config(n_gpus, segments_per_gpu=n_seg,
                                  compressed_bytes_per_value=round(c, 4),
                                  compression_ratio=round(4.0 / c, 3)),
        "encode_gbs": enc_gbs, "decode_gbs": dec_gbs,
        "encode_ms": enc_ms / args.steps, "decode_ms": dec_ms / args.steps,
        "kernel_ms": {k: round(v, 4) for k, v in kt.items()},
        "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches),
        "native_so_loaded": native_so_loaded(),
        "clocks": clk, "check": check, "wall_s": t_wall,
    }
    if gather is not None:
        line["gather"] = gather
    if halves is not None:
        line["k_encode_halves"] = halves
    if checksums is not None:
        line["checksums"] = checksums
    if refflow is not None:
        line["e2e_reference_flow"] = refflow
    if sweep is not None:
        line["sweep"] = sweep
    print(json.dumps(line), flush=True)
    if shard is not None:
        shard.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


def run_gather(torch, dist, shard, d_out, d_sizes, d_base, dev, rank):
    """Gather of the packed stream to rank 0 over NVLink (reported
    on its own, it is not part of the encode/decode metric): mnwnccl_GatherStream,
    one grouped ncclSend/ncclRecv at the prefix offsets. Device-timed."""
    sizes_h = [int(v) for v in d_sizes.cpu()]
    total = sum(sizes_h)
    prefix_ok = int(d_base.item()) == sum(sizes_h[:rank])
    free = torch.cuda.mem_get_info(dev)[0]
    ok = torch.tensor([0.0 if (rank == 0 and total > free - (2 << 30)) else 1.0], device=dev)
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    if float(ok.item()) == 0.0:
        return {"skipped": "sink has no room for the whole stream", "bytes": total}
    d_global = torch.empty(total, dtype=torch.uint8, device=dev) if rank == 0 else None
    shard.gather(d_out, sizes_h, 0, d_global)  # warm-up
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    shard.gather(d_out, sizes_h, 0, d_global)
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    # the sink's copy of this rank's slab must be the slab (checked on rank 0 for its own, cheaply)
    same = True
    if rank == 0:
        same = bool(torch.equal(d_global[:sizes_h[0]], d_out[:sizes_h[0]]))
    del d_global
    moved = total - sizes_h[0]
    gbs = moved / (float(ms.item()) * 1e-3) / 1e9
    return {"ms": float(ms.item()), "bytes_into_sink": moved, "gbs": gbs, "frac_of_900_gbs_ingress": gbs / 900.0,
            # what plain device-to-device copies from 1..3 peers into one GPU reach on this pool's boxes
            # (scripts/p2p_peak.py, measured once: 766 / 773 / 774.5 GB/s): the practical ingress ceiling
            "p2p_copy_ceiling_gbs": P2P_COPY_CEILING_GBS, "frac_of_p2p_copy_ceiling": gbs / P2P_COPY_CEILING_GBS,
            "prefix_ok": bool(prefix_ok), "sink_slab_ok": same,
            "note": "mnwnccl_GatherStream: one ncclGroupStart..ncclSend/ncclRecv x ranks..ncclGroupEnd into rank 0; "
                    "not part of the metric"}


def roundtrip_check(torch, d_in, d_dec, fields, seg):
    """|x - decode(encode(x))| <= maxDiff/2^depth < Delta (mod L for positions) on every value."""
    worst_pos = worst_vel = 0.0
    step = max(1, len(fields) // 16) | 1  # odd stride: visits positions and velocities
    for i in range(0, len(fields), step):
        off, n, code, delta, width, _ = fields[i]
        a = d_in[off:off + 3 * n].double()
        b = d_dec[off:off + 3 * n].double()
        err = (a - b).abs()
        if width > 0:
            err = torch.minimum(err, width - err)
            worst_pos = max(worst_pos, float(err.max()) / delta)
        else:
            worst_vel = max(worst_vel, float(err.max()) / delta)
    ok = worst_pos <= 1.0 + 1e-3 and worst_vel <= 1.0 + 1e-3
    if not ok:
        raise SystemExit("round-trip error bound violated: pos %.4f vel %.4f of Delta" % (worst_pos, worst_vel))
    return {"roundtrip_max_err_over_delta": {"pos": round(worst_pos, 4), "vel": round(worst_vel, 4)}, "ok": ok}


def run_e2e(torch, mc, codec, d_in, args, n_gpus, dev, world):
    """Same step through mnwcuda_EncodeFieldsHost / mnwcuda_DecodeFieldsHost on
    pinned host buffers: every step copies its input floats H2D, the payload
    D2H, the payload H2D again and the decoded floats D2H."""
    import torch.distributed as dist
    n_seg = min(args.e2e_segments, args.segments)
    seg = SEG
    fields = field_list(n_seg, seg, same_seed=args.same_seed)
    nf = len(fields)
    descs = mc.make_descs(fields)
    floats = nf * 3 * seg
    h_in = torch.empty(floats, dtype=torch.float32, pin_memory=True)
    h_in.copy_(d_in[:floats])
    h_dec = torch.empty(floats, dtype=torch.float32, pin_memory=True)
    cap = nf * mc.field_bound(seg)
    h_out = torch.empty(cap, dtype=torch.uint8, pin_memory=True)
    h_off = torch.zeros(nf + 1, dtype=torch.int64, pin_memory=True)
    torch.cuda.synchronize()
    steps = max(2, min(args.steps, 5))
    total = 0
    for _ in range(2):
        total = codec.encode_host(h_in, descs, nf, h_out, h_off)
        codec.decode_host(h_out, h_off, descs, nf, h_dec)
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    te = td = 0.0
    for _ in range(steps):
        a = time.perf_counter()
        total = codec.encode_host(h_in, descs, nf, h_out, h_off)
        b = time.perf_counter()
        codec.decode_host(h_out, h_off, descs, nf, h_dec)
        te += b - a
        td += time.perf_counter() - b
    elapsed = time.perf_counter() - t0
    t = torch.tensor([elapsed, te, td], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    elapsed, te, td = [float(v) for v in t.cpu()]
    raw = floats * 4
    # the ceiling: the same bytes over the same pinned buffers with no kernels, all ranks at once
    if world > 1:
        dist.barrier()
    tc = torch.tensor([host_ceiling(torch, h_in, h_out, dev, raw, int(total))], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tc, op=dist.ReduceOp.MAX)
    ceiling = n_gpus * 2 * raw / float(tc.item()) / 1e9
    value = n_gpus * 2 * raw * steps / elapsed / 1e9
    return {
        "value": value, "unit": UNIT,
        "h2d_bytes_per_step": int(raw + total), "d2h_bytes_per_step": int(total + raw),
        "encode_gbs": n_gpus * raw * steps / te / 1e9, "decode_gbs": n_gpus * raw * steps / td / 1e9,
        "host_ceiling_gbs": ceiling, "frac_of_ceiling": value / ceiling,
        "host_ceiling_note": "copies only (H2D raw || D2H packed, then H2D packed || D2H raw), same pinned buffers, "
                             "all ranks concurrently, max over ranks",
        "steps": steps,
        "sample": "%d of the %d segments per GPU (%.0f MiB raw), pinned host buffers, wall clock around the "
                  "synchronous C-ABI calls" % (n_seg, args.segments, raw / 2**20),
    }


def run_cpu_leg(args, d_in, fields, gpu_payload=None, gpu_fields=0):
    """The reference's CPU code timed on a bounded sample of the same workload: the
    slab's first segments, the device generator's floats copied to the host.
    gpu_payload(i) -> bytes of field i as the GPU arm encoded it; the two must be
    byte-identical (parity check in the same run)."""
    threads = os.cpu_count() or 1
    n_seg = args.cpu_segments or max(1, min(threads, 32))
    n_seg = min(n_seg, args.segments)
    threads = min(threads, n_seg)
    orc, kind = load_oracle()
    fl = d_in[:n_seg * 6 * SEG].cpu().numpy().reshape(n_seg, 2, 3 * SEG)
    pos = np.ascontiguousarray(fl[:, 0, :]).reshape(-1)
    vel = np.ascontiguousarray(fl[:, 1, :]).reshape(-1)
    del fl
    raw, payloads, one_step = cpu_baseline(orc, pos, vel, threads, n_seg, SEG, seeds=[f[5] for f in fields])
    parity = None
    if gpu_payload is not None:
        k = min(len(payloads), gpu_fields)
        same = sum(1 for i in range(k) if np.array_equal(gpu_payload(i), payloads[i]))
        parity = {"fields_compared": k, "byte_identical": same, "ok": same == k, "against": kind}
    one_step()
    te = td = 0.0
    reps = 0
    t0 = time.perf_counter()
    while reps < 2 or (time.perf_counter() - t0 < 12.0 and reps < 40):
        a, b = one_step()
        te += a
        td += b
        reps += 1
    out = {
        "value": 2 * raw * reps / (te + td) / 1e9, "unit": UNIT, "cores": threads, "kind": kind,
        "encode_gbs": raw * reps / te / 1e9, "decode_gbs": raw * reps / td / 1e9,
        "sample": "%d segments x 2^%d particles pos+vel (%.0f MiB raw) x %d passes, threads over fields, "
                  "device generator's floats" % (n_seg, SEG_LOG2, raw / 2**20, reps),
        "parity": parity,
    }
    out.update(extra_cpu_figures(orc, one_step, SEG))
    return out


def sharded_parity(torch, dist, d_in, d_out, d_off, fields, dev, pick=(0, 1)):
    """N > 1: every rank byte-compares some of its fields (floats of the 2048^3 box,
    copied to the host) with the CPU oracle; the verdict is all-reduced."""
    orc, kind = load_oracle()
    offs = d_off.cpu().numpy()
    same = 0
    for i in pick:
        off, n, code, delta, width, seed = fields[i]
        host = d_in[off:off + 3 * n].cpu().numpy()
        want = orc.encode_field(code, host, delta, width, seed=seed)
        got = d_out[int(offs[i]):int(offs[i + 1])].cpu().numpy()
        same += int(np.array_equal(got, want))
    t = torch.tensor([same, len(pick)], dtype=torch.int64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    same, total = [int(v) for v in t.cpu()]
    return {"fields_compared": total, "byte_identical": same, "ok": same == total, "against": kind,
            "note": "fields %s of every rank's slab" % (list(pick),)}


def read_headers(torch, d_out, d_off, nf):
    """The fields' 64-byte headers as a structured array (one gather on the device)."""
    offs = d_off[:nf].to(torch.int64)
    idx = offs[:, None] + torch.arange(64, device=d_out.device, dtype=torch.int64)[None, :]
    raw = d_out[idx.reshape(-1)].cpu().numpy().tobytes()
    hd = np.dtype([("Magic", "<u4"), ("Version", "<u4"), ("FieldCode", "<u4"), ("ParticleLen", "<i4"),
                   ("Depth", "u1"), ("Pad", "u1", 3), ("Width", "<f4"), ("X0", "<f4", 3), ("X1", "<f4", 3),
                   ("BlockLen", "<u4"), ("DitherRun", "<u2"), ("DitherLanes", "<u2"), ("DitherSeed", "<u8")])
    return np.frombuffer(raw, hd)


SWEEP_DEPTHS = (6, 8, 10, 12, 14, 16, 18, 20, 22, 24)


def run_sweep(torch, mc, codec, d_in, d_out, d_off, d_dec, fields, seg, peak, ngrid, first_id, n_seg):
    """BASELINE configs[4]: the same 1024^3 slab at quantisation depths 6..24. For
    every field Delta = maxDiff / 2^d * 1.0001, so that deltaToDepth
    (/root/reference/src/quant.c:675-697) returns exactly d; encode
    once, then time decode only (and the encoder at the same depth). Run on the
    smooth field of the headline and on a rough one (white-noise displacement of
    0.3 cells on top: deltas that do not shrink with the grid spacing)."""
    nf = len(fields)
    values = nf * 3 * seg
    out = {"depths": list(SWEEP_DEPTHS), "delta_rule": "maxDiff/2^d*1.0001 per field", "reps": 3, "fields": {}}

    def one_field(kind):
        codec.encode_dev(d_in, mc.make_descs(fields), nf, d_out, d_off)
        torch.cuda.synchronize()
        if codec.status() != 0:
            raise SystemExit("sweep: device-side error in the default encode")
        hd = read_headers(torch, d_out, d_off, nf)
        md = np.max(hd["X1"] - hd["X0"], axis=1).astype(np.float32)
        rows = []
        codec.enable_kernel_timing(True)
        for d in SWEEP_DEPTHS:
            deltas = (md / np.float32(2.0 ** d) * np.float32(1.0001)).astype(np.float32)
            fl = [(f[0], f[1], f[2], float(deltas[i]), f[4], f[5]) for i, f in enumerate(fields)]
            descs = mc.make_descs(fl)
            enc, dec = [], []
            for rep in range(4):  # first pass untimed
                codec.encode_dev(d_in, descs, nf, d_out, d_off)
                codec.decode_dev(d_out, d_off, descs, nf, d_dec)
                torch.cuda.synchronize()
                if rep:
                    kt = codec.kernel_times()
                    enc.append(kt["k_encode"])
                    dec.append(kt["k_unpack_dequant"])
            if codec.status() != 0:
                raise SystemExit("sweep: device-side error at depth %d" % d)
            hd2 = read_headers(torch, d_out, d_off, nf)
            depth_ok = bool((hd2["Depth"] == d).all())
            comp = int(d_off[nf].item())
            c = comp / values
            # error bound on two fields: |x - x'| <= Delta (mod L)
            # (the bound is Delta plus the float spacing at the values themselves: at depth
            # 24 a bin of a 100-wide box is narrower than one ulp of 100)
            worst = 0.0
            for i in (0, 1):
                off, n, code, delta, width, _ = fl[i]
                x = d_in[off:off + 3 * n].double()
                err = (x - d_dec[off:off + 3 * n].double()).abs()
                if width > 0:
                    err = torch.minimum(err, width - err)
                tol = delta + 2.0 * 2.0 ** -23 * max(float(x.abs().max()), width)
                worst = max(worst, float(err.max()) / tol)
            em, dm = float(np.mean(enc)), float(np.mean(dec))
            alg = (4.0 + c) * values
            rows.append({
                "depth": d, "depth_ok": depth_ok, "compressed_bytes_per_value": round(c, 4),
                "payload_bits_per_value": round(8.0 * (comp - nf * mc.field_fixed_bytes(seg)) / values, 3),
                "decode_ms": round(dm, 4), "decode_raw_gbs": round(4.0 * values / (dm * 1e-3) / 1e9, 1),
                "decode_achieved_gbs": round(alg / (dm * 1e-3) / 1e9, 1), "decode_frac": round(alg / (dm * 1e-3) / 1e9 / peak, 4),
                "encode_ms": round(em, 4), "encode_achieved_gbs": round(alg / (em * 1e-3) / 1e9, 1),
                "encode_frac": round(alg / (em * 1e-3) / 1e9 / peak, 4),
                "roundtrip_max_err_over_tol": round(worst, 4),
                "ok": bool(depth_ok and worst <= 1.0 + 1e-3),
            })
        codec.enable_kernel_timing(False)
        out["fields"][kind] = rows

    one_field("smooth")
    generate(codec, d_in, ngrid, first_id, n_seg, seg, noise_cells=0.3)
    torch.cuda.synchronize()
    # the rough field at the headline's own accuracies (pos 1e-3, vel 1), then the sweep
    descs = mc.make_descs(fields)
    codec.enable_kernel_timing(True)
    enc, dec = [], []
    for rep in range(4):
        codec.encode_dev(d_in, descs, nf, d_out, d_off)
        codec.decode_dev(d_out, d_off, descs, nf, d_dec)
        torch.cuda.synchronize()
        if rep:
            kt = codec.kernel_times()
            enc.append(kt["k_encode"])
            dec.append(kt["k_unpack_dequant"])
    codec.enable_kernel_timing(False)
    comp = int(d_off[nf].item())
    c = comp / values
    alg = (4.0 + c) * values
    out["rough_default_accuracy"] = {
        "noise_cells": 0.3, "compressed_bytes_per_value": round(c, 4),
        "payload_bits_per_value": round(8.0 * (comp - nf * mc.field_fixed_bytes(seg)) / values, 3),
        "encode_ms": round(float(np.mean(enc)), 4), "encode_frac": round(alg / (np.mean(enc) * 1e-3) / 1e9 / peak, 4),
        "decode_ms": round(float(np.mean(dec)), 4), "decode_frac": round(alg / (np.mean(dec) * 1e-3) / 1e9 / peak, 4),
    }
    one_field("rough")
    out["ok"] = all(r["ok"] for rows in out["fields"].values() for r in rows)
    generate(codec, d_in, ngrid, first_id, n_seg, seg)  # back to the headline's field
    torch.cuda.synchronize()
    return out


def run_encode_halves(torch, codec, d_in, descs, nf, d_out, d_off, alg_bytes, peak):
    """k_encode's two halves on their own (explains the fused kernel's time; untimed in
    the metric, and the output of these launches is not a valid stream): the B work
    alone -- quantise .. pack .. place with the parameters the previous full launch left
    behind, i.e. the one-pass encoder for callers that know the bounds, every read from
    DRAM -- and the A work alone (the streaming min/max through the same ring)."""
    codec.enable_kernel_timing(True)

    def timed(reps=3):
        ms = []
        for _ in range(reps):
            codec.encode_dev(d_in, descs, nf, d_out, d_off)
            torch.cuda.synchronize()
            ms.append(codec.kernel_times()["k_encode"])
        return float(np.mean(ms))

    fused = timed()
    os.environ["MNWCUDA_ENCODE_ONLY"] = "B"
    b_only = timed()
    del os.environ["MNWCUDA_ENCODE_ONLY"]
    old_dbg = os.environ.get("MNWCUDA_DBG")
    os.environ["MNWCUDA_DBG"] = "64"
    a_only = timed()
    if old_dbg is None:
        del os.environ["MNWCUDA_DBG"]
    else:
        os.environ["MNWCUDA_DBG"] = old_dbg
    codec.status()  # (clears whatever the A-only launches raised: their output is never finalised properly)
    codec.encode_dev(d_in, descs, nf, d_out, d_off)  # leave a valid stream behind for the legs that follow
    torch.cuda.synchronize()
    codec.enable_kernel_timing(False)
    if codec.status() != 0:
        raise SystemExit("device-side error after the encode-halves experiment")
    raw = nf * 3 * SEG * 4
    return {"fused_ms": round(fused, 4), "b_only_ms": round(b_only, 4), "a_only_ms": round(a_only, 4),
            "b_only_frac": round(alg_bytes / (b_only * 1e-3) / 1e9 / peak, 4),
            "a_only_read_gbs": round(raw / (a_only * 1e-3) / 1e9, 1),
            "a_only_frac_of_peak": round(raw / (a_only * 1e-3) / 1e9 / peak, 4),
            "note": "B only = the one-pass encoder with given quantisation parameters (MNWCUDA_ENCODE_ONLY=B), "
                    "A only = min/max pass through the same TMA ring (MNWCUDA_DBG=64); timing experiments"}


def run_checksums(torch, mc, codec, d_out, d_off, descs, nf, comp_bytes, peak):
    """The segment image's BlockHeader table for the whole packed slab: util_Checksum
    (/root/reference/src/util.c:438-445) of every block on the device, one thread per
    block (untimed in the metric: the reference leaves checksums to the caller)."""
    codec.enable_kernel_timing(True)
    ms = []
    for rep in range(4):
        hdr = codec.block_headers_dev(d_out, d_off, descs, nf)
        torch.cuda.synchronize()
        if rep:
            ms.append(codec.kernel_times()["k_block_headers"])
    codec.enable_kernel_timing(False)
    if codec.status() != 0:
        raise SystemExit("device-side error in the block checksums")
    t = float(np.mean(ms))
    return {"kernel": "k_block_headers", "blocks": int(hdr.shape[0]), "ms": round(t, 4),
            "packed_gbs": round(comp_bytes / (t * 1e-3) / 1e9, 1), "frac_of_hbm_peak": round(comp_bytes / (t * 1e-3) / 1e9 / peak, 4),
            "note": "reads the packed stream once (algorithmic bytes = packed bytes), writes 8 B per block"}


def run_reference_flow(torch, mc, d_in, seg, fields):
    """One segment (positions, velocities, ids; 2^22 particles) through the host
    library the way /root/reference/src/funcs.c drives a plugin: (a) literally,
    quant_QField -> Register_GetCompressor -> CFunc -> checksum -> DFunc ->
    quant_Field with one Compressor / Decompressor per field (src/funcs.c:13-76,
    122-189; uint64 QFields cross the host boundary), (b) fused, MnwSeg_Encode /
    MnwSeg_Decode. Wall clock, host arrays in and out."""
    import ctypes as C
    h = mc.host_lib()
    pos = d_in[fields[0][0]:fields[0][0] + 3 * seg].cpu().numpy()
    vel = d_in[fields[1][0]:fields[1][0] + 3 * seg].cpu().numpy()
    ids = np.arange(seg, dtype=np.uint64)
    pacc = mc.PositionAccuracy(None, POS_DELTA, BOX, 0)
    vacc = mc.VelocityAccuracy(None, VEL_DELTA, 0, 0, 0.0)
    iacc = mc.IDAccuracy(1024)
    fl = (mc.Field * 3)()
    fl[0] = mc.Field(mc.FieldHeader(FIELD_POSN, mc.ALGO_DIFF, mc.ALGO_VERSION, seg), 1, pos.ctypes.data, C.addressof(pacc))
    fl[1] = mc.Field(mc.FieldHeader(FIELD_VELC, mc.ALGO_DIFF, mc.ALGO_VERSION, seg), 1, vel.ctypes.data, C.addressof(vacc))
    fl[2] = mc.Field(mc.FieldHeader(mc.FIELD_PTID, mc.ALGO_DIFF, mc.ALGO_VERSION, seg), 1, ids.ctypes.data, C.addressof(iacc))
    raw = pos.nbytes + vel.nbytes + ids.nbytes
    cfunc = C.CFUNCTYPE(mc.CField, mc.QField, C.c_void_p)
    dfunc = C.CFUNCTYPE(mc.QField, mc.CField, C.c_void_p)

    def literal():
        reg = h.Register_New()
        cfs, ok = [], True
        for i in range(3):
            qf = h.quant_QField(fl[i])
            comp = h.Register_GetCompressor(reg, mc.ALGO_DIFF, mc.ALGO_VERSION)
            cf = cfunc(comp.CFunc)(qf, comp.Buffer)
            cf.Checksum = h.MnwChecksum(cf.Data, cf.DataLen)
            h.Register_FreeCompressor(reg, mc.ALGO_DIFF, mc.ALGO_VERSION, comp)
            h.quant_FreeQField(qf)
            cfs.append(cf)
        for i in range(3):
            dec = h.Register_GetDecompressor(reg, mc.ALGO_DIFF, mc.ALGO_VERSION)
            qf = dfunc(dec.DFunc)(cfs[i], dec.Buffer)
            f = h.quant_Field(qf)
            if i == 2:
                back = np.ctypeslib.as_array(C.cast(f.Data, C.POINTER(C.c_uint64)), shape=(seg,))
                ok = bool(np.array_equal(back, ids))
            h.Register_FreeDecompressor(reg, mc.ALGO_DIFF, mc.ALGO_VERSION, dec)
            h.quant_FreeQField(qf)
            h.quant_FreeField(f)
            mc._libc.free(cfs[i].Data)
        h.Register_Free(reg)
        return ok

    def fused():
        cs = h.MnwSeg_Encode(mc.Seg(fl, 3))
        s = h.MnwSeg_Decode(cs)
        ok = all(s.Fields[i].Valid == 1 for i in range(3))
        h.MnwSeg_Free(s)
        h.MnwCSeg_Free(cs)
        return ok

    out = {"segment": "2^%d particles: positions, velocities, ids (%.0f MiB raw)" % (SEG_LOG2, raw / 2**20)}
    for name, fn in (("literal", literal), ("fused", fused)):
        ok = fn()
        t0 = time.perf_counter()
        reps = 3
        for _ in range(reps):
            ok = fn() and ok
        el = (time.perf_counter() - t0) / reps
        out[name] = {"ms": round(1e3 * el, 2), "gbs": round(2 * raw / el / 1e9, 3), "ok": bool(ok)}
    out["note"] = ("encode + decode of one segment, wall clock; literal = uint64 QFields over the host boundary "
                   "(8 B per value each way), one pooled context per field")
    return out


def host_ceiling(torch, h_in, h_out_bytes, dev, raw, comp, steps=3):
    """What pinned host memory and PCIe allow for the e2e copy pattern with no
    kernels at all: (H2D raw || D2H compressed) then (H2D compressed || D2H raw),
    plain cudaMemcpyAsync on two streams."""
    d_raw = torch.empty(raw, dtype=torch.uint8, device=dev)
    d_comp = torch.empty(comp, dtype=torch.uint8, device=dev)
    h_raw = h_in.view(torch.uint8)[:raw]
    h_comp = h_out_bytes[:comp]
    s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

    def phase(a_dst, a_src, b_dst, b_src):
        with torch.cuda.stream(s1):
            a_dst.copy_(a_src, non_blocking=True)
        with torch.cuda.stream(s2):
            b_dst.copy_(b_src, non_blocking=True)
        s1.synchronize()
        s2.synchronize()

    phase(d_raw, h_raw, h_comp, d_comp)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        phase(d_raw, h_raw, h_comp, d_comp)   # encode's copies
        phase(d_comp, h_comp, h_raw, d_raw)   # decode's copies
    el = time.perf_counter() - t0
    del d_raw, d_comp
    return el / steps


def load_traffic(kernel):
    """dram bytes per launch of `kernel` from the committed ncu --set full summary, or None."""
    path = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        with open(path) as fh:
            return json.load(fh).get(kernel)
    except Exception:
        return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--segments", type=int, default=256, help="segments of 2^22 particles per GPU")
    ap.add_argument("--e2e-segments", type=int, default=32)
    ap.add_argument("--cpu-segments", type=int, default=0, help="CPU arm sample (0 = from the core count)")
    ap.add_argument("--encode-only", action="store_true", help="kernel experiments: time encode only, no checks")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer e2e leg (kernel experiments)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-gather", action="store_true", help="N>1: skip the (separately reported) stream gather")
    ap.add_argument("--no-sweep", action="store_true", help="skip the depth sweep (BASELINE configs[4])")
    ap.add_argument("--same-seed", action="store_true",
                    help="one dither seed for all fields (one shared stream table) instead of one per field")
    args = ap.parse_args()
    if args.steps < 1:
        args.steps = 1
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
