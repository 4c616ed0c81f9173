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
def field_list(n_segments, seg=SEG):
    """[(float offset, n, code, delta, width, seed)]: field 2s = positions of
    segment s, 2s+1 = velocities; each 3*seg planar floats."""
    import minnow_c_b200 as mc
    out = []
    for s in range(n_segments):
        out.append(((2 * s) * 3 * seg, seg, mc.FIELD_POSN, POS_DELTA, BOX, 0))
        out.append(((2 * s + 1) * 3 * seg, seg, mc.FIELD_VELC, VEL_DELTA, 0.0, 0))
    return out


def generate(codec, d_in, ngrid, first_id, n_segments, seg=SEG):
    for s in range(n_segments):
        pos = d_in[(2 * s) * 3 * seg:(2 * s + 1) * 3 * seg]
        vel = d_in[(2 * s + 1) * 3 * seg:(2 * s + 2) * 3 * seg]
        codec.generate(ngrid, BOX, GEN_SEED, first_id + s * seg, seg, pos, vel)


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


def host_sample(n_fields_half, seg, ngrid):
    """n_fields_half segments of pos+vel as host arrays. Uses the device
    generator when a GPU is present (the same floats as the GPU arm), else the
    numpy twin of tests/_data.py."""
    try:
        import torch
        have_gpu = torch.cuda.is_available()
    except Exception:
        have_gpu = False
    if have_gpu:
        import torch
        import minnow_c_b200 as mc
        codec = mc.Codec(0)
        d = torch.empty(n_fields_half * 6 * seg, dtype=torch.float32, device="cuda:0")
        generate(codec, d, ngrid, 0, n_fields_half, seg)
        torch.cuda.synchronize()
        h = d.cpu().numpy()
        codec.close()
        del d
        return h, "device generator"
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from _data import zeldovich_host
    chunks = []
    for s in range(n_fields_half):
        pos, vel = zeldovich_host(ngrid, L=BOX, seed=GEN_SEED % 65536, start=s * seg, count=seg)
        chunks += [pos.reshape(-1), (2.0 * vel).reshape(-1)]
    return np.concatenate(chunks), "numpy twin"


def cpu_baseline(threads, n_seg, seg, ngrid):
    orc, kind = load_oracle()
    FIELD_POSN, FIELD_VELC = 0x506F736E, 0x56656C63
    h, gen = host_sample(n_seg, seg, ngrid)
    fl = h.reshape(n_seg, 2, 3 * seg)
    pos = np.ascontiguousarray(fl[:, 0, :]).reshape(-1)
    vel = np.ascontiguousarray(fl[:, 1, :]).reshape(-1)
    del h, fl
    payloads = []
    for s in range(n_seg):
        payloads.append(orc.encode_field(FIELD_POSN, pos[s * 3 * seg:(s + 1) * 3 * seg], POS_DELTA, BOX))
        payloads.append(orc.encode_field(FIELD_VELC, vel[s * 3 * seg:(s + 1) * 3 * seg], VEL_DELTA, 0.0))
    raw = 2 * n_seg * 3 * seg * 4

    def one_step():
        te = 0.0
        for code, data, delta, width in ((FIELD_POSN, pos, POS_DELTA, BOX), (FIELD_VELC, vel, VEL_DELTA, 0.0)):
            secs, _ = orc.bench_encode(code, data, seg, n_seg, delta, width, threads)
            te += secs
        td = orc.bench_decode(payloads, seg, threads)
        return te, td

    return orc, kind, gen, raw, payloads, one_step


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    threads = os.cpu_count() or 1
    n_seg = args.cpu_segments or max(1, min(threads, 32))
    threads = min(threads, n_seg)
    ngrid = 1024 if args.gpus == 1 else 2048
    orc, kind, gen, raw, _payloads, one_step = cpu_baseline(threads, n_seg, SEG, ngrid)
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
    sample = ("%d segments x 2^%d particles, pos+vel (%.0f MiB raw) per step, %s; encode %.3f GB/s, decode %.3f GB/s"
              % (n_seg, SEG_LOG2, raw / 2**20, gen, raw * args.steps / te / 1e9, raw * args.steps / td / 1e9))
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * (te + td) / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32/u32",
        "data": "synthetic",
        "config": workload_config(args.gpus, sample_only=True),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "wall_s": wall,
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
    fields = field_list(n_seg, seg)
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

    def step(record=None):
        if record:
            record[0].record()
        codec.encode_dev(d_in, descs, nf, d_out, d_off)
        if world > 1:
            # output placement of the sharded stream: all-gather of per-rank
            # compressed sizes, exclusive prefix = this rank's base offset
            _sizes, _base = sharded.exchange_sizes(d_off[nf:nf + 1])
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
        gather = run_gather(torch, dist, sharded, d_out, d_off, nf, dev, rank)

    value = n_gpus * 2 * raw_bytes * args.steps / (tot_ms * 1e-3) / 1e9
    enc_gbs = n_gpus * raw_bytes * args.steps / (enc_ms * 1e-3) / 1e9
    dec_gbs = n_gpus * raw_bytes * args.steps / (dec_ms * 1e-3) / 1e9

    if rank != 0:
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
        "alg_bytes_per_value": 4.0 + c, "values_per_launch": nf * 3 * seg,
        "kernels": {k: {"ms": round(v["ms"], 4), "achieved_gbs": round(v["achieved"], 1),
                        "frac": round(v["achieved"] / peak, 4)} for k, v in roof.items()},
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

        cpu = run_cpu_leg(args, gpu_payload, nf)
        if cpu.get("parity") is not None:
            check["sample_bytes_vs_reference"] = cpu["parity"]
            check["ok"] = bool(check["ok"] and cpu["parity"]["ok"])

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
        "clocks": clk, "check": check, "wall_s": t_wall,
    }
    if gather is not None:
        line["gather"] = gather
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


def run_gather(torch, dist, sharded, d_out, d_off, nf, dev, rank):
    """Gather of the packed stream to rank 0 over NVLink (reported
    on its own, it is not part of the encode/decode metric). Device-timed."""
    sizes, _ = sharded.exchange_sizes(d_off[nf:nf + 1])
    sizes_h = [int(v) for v in sizes.cpu()]
    total = sum(sizes_h)
    free = torch.cuda.mem_get_info(dev)[0]
    if rank == 0 and total > free - (2 << 30):
        ok = torch.zeros(1, device=dev)
    else:
        ok = torch.ones(1, device=dev)
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    if float(ok.item()) == 0.0:
        return {"skipped": "sink has no room for the whole stream", "bytes": total}
    local = d_out[:sizes_h[rank]]
    sharded.gather_stream(local, sizes_h, dst=0)  # warm-up
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    out = sharded.gather_stream(local, sizes_h, dst=0)
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    del out
    moved = total - sizes_h[0]
    return {"ms": float(ms.item()), "bytes_into_sink": moved, "gbs": moved / (float(ms.item()) * 1e-3) / 1e9,
            "note": "ncclSend/Recv of each rank's packed slab to rank 0; not part of the metric"}


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
    fields = field_list(n_seg, seg)
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
    return {
        "value": n_gpus * 2 * raw * steps / elapsed / 1e9, "unit": UNIT,
        "h2d_bytes_per_step": int(raw + total), "d2h_bytes_per_step": int(total + raw),
        "encode_gbs": n_gpus * raw * steps / te / 1e9, "decode_gbs": n_gpus * raw * steps / td / 1e9,
        "steps": steps,
        "sample": "%d of the %d segments per GPU (%.0f MiB raw), pinned host buffers, wall clock around the "
                  "synchronous C-ABI calls" % (n_seg, args.segments, raw / 2**20),
    }


def run_cpu_leg(args, gpu_payload=None, gpu_fields=0):
    """The reference's CPU code timed on a bounded sample of the same workload.
    gpu_payload(i) -> bytes of field i as the GPU arm encoded it: the sample is
    the slab's first segments, so the two must be byte-identical (SURVEY 8d:
    parity check in the same run)."""
    threads = os.cpu_count() or 1
    n_seg = args.cpu_segments or max(1, min(threads, 32))
    threads = min(threads, n_seg)
    orc, kind, gen, raw, payloads, one_step = cpu_baseline(threads, n_seg, SEG, 1024)
    parity = None
    if gpu_payload is not None and gen == "device generator":
        k = min(len(payloads), gpu_fields)
        same = sum(1 for i in range(k) if np.array_equal(gpu_payload(i), payloads[i]))
        parity = {"fields_compared": k, "byte_identical": same, "ok": same == k}
    one_step()
    te = td = 0.0
    reps = 0
    t0 = time.perf_counter()
    while reps < 2 or (time.perf_counter() - t0 < 12.0 and reps < 40):
        a, b = one_step()
        te += a
        td += b
        reps += 1
    return {
        "value": 2 * raw * reps / (te + td) / 1e9, "unit": UNIT, "cores": threads, "kind": kind,
        "encode_gbs": raw * reps / te / 1e9, "decode_gbs": raw * reps / td / 1e9,
        "sample": "%d segments x 2^%d particles pos+vel (%.0f MiB raw) x %d passes, threads over fields, %s"
                  % (n_seg, SEG_LOG2, raw / 2**20, reps, gen),
        "parity": parity,
    }


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
    args = ap.parse_args()
    if args.steps < 1:
        args.steps = 1
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
