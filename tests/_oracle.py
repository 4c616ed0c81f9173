"""ctypes front-end for the test-only CPU oracle (oracle/liboracle.so and,
when built, oracle/_ref/liboracle_ref.so + oracle/_ref/libminnow_ref.so).

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs import
this module. The product package never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")

FIELD_POSN = 0x506F736E
FIELD_VELC = 0x56656C63
FIELD_PTID = 0x50746964
FIELD_UNSI = 0x556E7369
WIDTH_DAMAGED = 0xFFFFFFFF
ALGO_DIFF = 0x44696666
BLOCK_LEN = 1024

_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_u32p = np.ctypeslib.ndpointer(np.uint32, flags="C_CONTIGUOUS")
_u64p = np.ctypeslib.ndpointer(np.uint64, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")

HEADER_DTYPE = np.dtype([
    ("Magic", "<u4"), ("Version", "<u4"), ("FieldCode", "<u4"),
    ("ParticleLen", "<i4"), ("Depth", "u1"), ("Pad", "u1", 3),
    ("Width", "<f4"), ("X0", "<f4", 3), ("X1", "<f4", 3),
    ("BlockLen", "<u4"), ("DitherRun", "<u2"), ("DitherLanes", "<u2"),
    ("DitherSeed", "<u8"),
])
assert HEADER_DTYPE.itemsize == 64
ENTRY_DTYPE = np.dtype([("First", "<u4"), ("MinDelta", "<i4"),
                        ("Width", "<u4"), ("Offset16", "<u4")])


def build(force=False):
    """Compile liboracle.so (and oracle/_ref when /root/reference exists)."""
    so = os.path.join(ORACLE_DIR, "liboracle.so")
    src = os.path.join(ORACLE_DIR, "minnow_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "all"],
                              stdout=subprocess.DEVNULL)
    return so


def _bind(lib):
    sig = {
        "orc_undo_periodic": (None, [_f32p, C.c_int64, C.c_float]),
        "orc_periodic": (None, [_f32p, C.c_int64, C.c_float]),
        "orc_minmax": (None, [_f32p, C.c_int64, C.POINTER(C.c_float), C.POINTER(C.c_float)]),
        "orc_delta_to_depth": (C.c_int, [C.c_float, C.c_float, C.c_float]),
        "orc_uniform_bin_index": (None, [_f32p, C.c_int64, C.c_uint8, C.c_float, C.c_float, _u32p]),
        "orc_undo_uniform_bin_index": (None, [_u32p, C.c_int64, C.c_uint8, C.c_float, C.c_float, _u64p, _f32p]),
        "orc_pack_words": (C.c_int64, [C.c_int64, C.c_uint8]),
        "orc_uniform_pack": (None, [_u32p, C.c_int64, C.c_uint8, _u32p]),
        "orc_undo_uniform_pack": (None, [_u32p, C.c_uint8, C.c_int64, _u32p]),
        "orc_checksum": (C.c_uint32, [_u8p, C.c_int64]),
        "orc_rand_seed": (None, [C.c_uint64, C.c_int64, _u64p]),
        "orc_rand_u64": (C.c_uint64, [_u64p]),
        "orc_rand_float": (C.c_float, [_u64p]),
        "orc_rand_jump": (None, [_u64p]),
        "orc_diff_nblocks": (C.c_int64, [C.c_int32]),
        "orc_diff_fixed_bytes": (C.c_int64, [C.c_int32]),
        "orc_diff_bound": (C.c_int64, [C.c_int32]),
        "orc_diff_nstreams": (C.c_int64, [C.c_int32]),
        "orc_diff_encode_field": (C.c_int64, [C.c_uint32, _f32p, C.c_int32, C.c_float, C.c_float,
                                               C.c_uint64, _u8p, C.c_int64, C.c_void_p]),
        "orc_diff_decode_field": (C.c_int, [_u8p, C.c_int64, _f32p, C.c_void_p]),
        "orc_diff_encode_q": (C.c_int64, [C.c_void_p, _u64p, _u8p, C.c_int64]),
        "orc_diff_decode_q": (C.c_int, [_u8p, C.c_int64, _u64p]),
        "orc_quantize_field": (C.c_int, [C.c_uint32, _f32p, C.c_int32, C.c_float, C.c_float,
                                          C.c_void_p, _u32p]),
        "orc_bench_encode": (C.c_double, [C.c_uint32, _f32p, C.c_int32, C.c_int64, C.c_float,
                                           C.c_float, C.c_int, C.POINTER(C.c_int64)]),
        "orc_bench_decode": (C.c_double, [C.POINTER(C.c_void_p), C.POINTER(C.c_int64), C.c_int32,
                                           C.c_int64, C.c_int]),
        "orc_dequantize_field": (C.c_int, [C.c_void_p, _u64p, _f32p]),
        "orc_id_quantize": (C.c_int, [_u64p, C.c_int32, C.c_uint64, _u64p, _u64p, _u64p]),
        "orc_id_dequantize": (C.c_int, [_u64p, C.c_int32, C.c_uint64, _u64p, _u64p]),
        "orc_uint_quantize": (C.c_int, [_u64p, C.c_int32, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), _u64p]),
        "orc_uint_dequantize": (C.c_int, [_u64p, C.c_int32, C.c_uint64, _u64p]),
        "orc_diff_encode_id_field": (C.c_int64, [_u64p, C.c_int32, C.c_uint64, _u8p, C.c_int64]),
        "orc_diff_decode_id_field": (C.c_int, [_u8p, C.c_int64, _u64p]),
        "orc_diff_encode_uint_field": (C.c_int64, [_u64p, C.c_int32, _u8p, C.c_int64]),
        "orc_diff_decode_uint_field": (C.c_int, [_u8p, C.c_int64, C.c_int32, _u64p]),
        "orc_seg_field_blocks": (C.c_int64, [C.c_void_p]),
        "orc_seg_bytes_bound": (C.c_int64, [C.c_void_p, C.c_int32]),
        "orc_seg_to_bytes": (C.c_int64, [C.c_void_p, C.c_int32, C.c_int32, _u8p, C.c_int64]),
        "orc_seg_from_bytes": (C.c_int32, [_u8p, C.c_int64, C.c_void_p, C.c_int32, C.POINTER(C.c_int32),
                                            C.POINTER(C.c_int32)]),
        "orc_gen_zeldovich": (C.c_int, [C.c_int32, C.c_float, C.c_uint64, C.c_int64, C.c_int64, C.c_float,
                                        C.c_void_p, C.c_void_p, C.c_int]),
        "orc_flavour": (C.c_char_p, []),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    return lib


class SegField(C.Structure):  # OrcSegField
    _fields_ = [("FieldCode", C.c_uint32), ("AlgoCode", C.c_uint32), ("AlgoVersion", C.c_uint32),
                ("ParticleLen", C.c_int32), ("Data", C.c_void_p), ("DataLen", C.c_int64)]


class Oracle:
    """Thin numpy wrapper around one build of the oracle."""

    def __init__(self, path):
        self.path = path
        self.lib = _bind(C.CDLL(path))
        self.flavour = self.lib.orc_flavour().decode()

    # ---- primitives ----
    def undo_periodic(self, x, L):
        x = np.array(x, dtype=np.float32)
        self.lib.orc_undo_periodic(x, x.size, L)
        return x

    def periodic(self, x, L):
        x = np.array(x, dtype=np.float32)
        self.lib.orc_periodic(x, x.size, L)
        return x

    def minmax(self, x):
        x = np.ascontiguousarray(x, dtype=np.float32)
        lo, hi = C.c_float(), C.c_float()
        self.lib.orc_minmax(x, x.size, C.byref(lo), C.byref(hi))
        return np.float32(lo.value), np.float32(hi.value)

    def delta_to_depth(self, delta, x0, x1):
        return self.lib.orc_delta_to_depth(delta, x0, x1)

    def uniform_bin_index(self, x, level, x0, dx):
        x = np.ascontiguousarray(x, dtype=np.float32)
        q = np.zeros(max(x.size, 1), np.uint32)
        self.lib.orc_uniform_bin_index(x, x.size, level, x0, dx, q)
        return q[:x.size]

    def undo_uniform_bin_index(self, q, level, x0, dx, state):
        q = np.ascontiguousarray(q, dtype=np.uint32)
        out = np.zeros(max(q.size, 1), np.float32)
        self.lib.orc_undo_uniform_bin_index(q, q.size, level, x0, dx, state, out)
        return out[:q.size]

    def pack(self, x, width):
        x = np.ascontiguousarray(x, dtype=np.uint32)
        nw = self.lib.orc_pack_words(x.size, width)
        out = np.zeros(max(nw, 1), np.uint32)
        self.lib.orc_uniform_pack(x if x.size else np.zeros(1, np.uint32), x.size, width, out)
        return out[:nw]

    def unpack(self, words, width, n):
        words = np.ascontiguousarray(words, dtype=np.uint32)
        if words.size == 0:
            words = np.zeros(1, np.uint32)
        out = np.zeros(max(n, 1), np.uint32)
        self.lib.orc_undo_uniform_pack(words, width, n, out)
        return out[:n]

    def checksum(self, data):
        data = np.ascontiguousarray(data, dtype=np.uint8)
        return self.lib.orc_checksum(data if data.size else np.zeros(1, np.uint8), data.size)

    def rand_seed(self, seed, n):
        st = np.zeros(2 * n, np.uint64)
        self.lib.orc_rand_seed(seed, n, st)
        return st.reshape(n, 2)

    # ---- composition ----
    def encode_field(self, field_code, data, delta, width=0.0, seed=0, want_q=False):
        data = np.ascontiguousarray(data, dtype=np.float32).reshape(-1)
        n = data.size // 3
        cap = self.lib.orc_diff_bound(n)
        out = np.zeros(cap, np.uint8)
        q = np.zeros(max(3 * n, 1), np.uint32) if want_q else None
        nbytes = self.lib.orc_diff_encode_field(
            field_code, data if data.size else np.zeros(1, np.float32), n, delta, width, seed,
            out, cap, q.ctypes.data if want_q else None)
        if nbytes < 0:
            raise ValueError("oracle encode failed: %d" % nbytes)
        return (out[:nbytes].copy(), q[:3 * n]) if want_q else out[:nbytes].copy()

    def decode_field(self, payload, want_q=False):
        payload = np.ascontiguousarray(payload, dtype=np.uint8)
        hd = parse_header(payload)
        n = int(hd["ParticleLen"])
        out = np.zeros(max(3 * n, 1), np.float32)
        q = np.zeros(max(3 * n, 1), np.uint32) if want_q else None
        rc = self.lib.orc_diff_decode_field(payload, payload.size, out,
                                            q.ctypes.data if want_q else None)
        if rc != 0:
            raise ValueError("oracle decode failed: %d" % rc)
        return (out[:3 * n], q[:3 * n]) if want_q else out[:3 * n]

    def quantize_field(self, field_code, data, delta, width=0.0):
        data = np.ascontiguousarray(data, dtype=np.float32).reshape(-1)
        n = data.size // 3
        hd = np.zeros(1, HEADER_DTYPE)
        q = np.zeros(max(3 * n, 1), np.uint32)
        rc = self.lib.orc_quantize_field(field_code, data, n, delta, width, hd.ctypes.data, q)
        if rc != 0:
            raise ValueError("oracle quantize failed: %d" % rc)
        return hd[0], q[:3 * n]

    def encode_q(self, hd, q64):
        hd = np.array([hd], dtype=HEADER_DTYPE)
        n = int(hd[0]["ParticleLen"])
        q64 = np.ascontiguousarray(q64, dtype=np.uint64)
        cap = self.lib.orc_diff_bound(n)
        out = np.zeros(cap, np.uint8)
        nbytes = self.lib.orc_diff_encode_q(hd.ctypes.data, q64, out, cap)
        if nbytes < 0:
            raise ValueError("oracle encode_q failed: %d" % nbytes)
        return out[:nbytes].copy()

    def decode_q(self, payload):
        payload = np.ascontiguousarray(payload, dtype=np.uint8)
        n = int(parse_header(payload)["ParticleLen"])
        q = np.zeros(max(3 * n, 1), np.uint64)
        rc = self.lib.orc_diff_decode_q(payload, payload.size, q)
        if rc != 0:
            raise ValueError("oracle decode_q failed: %d" % rc)
        return q[:3 * n]

    # ---- quantisers alone, integer fields, segment image ----
    def dequantize_field(self, hd, q64):
        hd = np.array([hd], dtype=HEADER_DTYPE)
        n = int(hd[0]["ParticleLen"])
        q64 = np.ascontiguousarray(q64, dtype=np.uint64)
        out = np.zeros(max(3 * n, 1), np.float32)
        rc = self.lib.orc_dequantize_field(hd.ctypes.data, q64 if q64.size else np.zeros(1, np.uint64), out)
        if rc:
            raise ValueError("orc_dequantize_field: %d" % rc)
        return out[:3 * n]

    def id_quantize(self, ids, width):
        ids = np.ascontiguousarray(ids, dtype=np.uint64)
        n = ids.size
        x0, x1, q = np.zeros(3, np.uint64), np.zeros(3, np.uint64), np.zeros(3 * n, np.uint64)
        rc = self.lib.orc_id_quantize(ids, n, width, x0, x1, q)
        if rc:
            raise ValueError("orc_id_quantize: %d" % rc)
        return x0, x1, q

    def id_dequantize(self, q, width, x0):
        q = np.ascontiguousarray(q, dtype=np.uint64)
        n = q.size // 3
        ids = np.zeros(n, np.uint64)
        rc = self.lib.orc_id_dequantize(q, n, width, np.ascontiguousarray(x0, dtype=np.uint64), ids)
        if rc:
            raise ValueError("orc_id_dequantize: %d" % rc)
        return ids

    def uint_quantize(self, data):
        data = np.ascontiguousarray(data, dtype=np.uint64)
        x0, x1 = C.c_uint64(), C.c_uint64()
        q = np.zeros(data.size, np.uint64)
        rc = self.lib.orc_uint_quantize(data, data.size, C.byref(x0), C.byref(x1), q)
        if rc:
            raise ValueError("orc_uint_quantize: %d" % rc)
        return x0.value, x1.value, q

    def encode_id_field(self, ids, width):
        ids = np.ascontiguousarray(ids, dtype=np.uint64)
        cap = self.lib.orc_diff_bound(ids.size)
        out = np.zeros(cap, np.uint8)
        nbytes = self.lib.orc_diff_encode_id_field(ids if ids.size else np.zeros(1, np.uint64), ids.size, width, out, cap)
        if nbytes < 0:
            raise ValueError("orc_diff_encode_id_field: %d" % nbytes)
        return out[:nbytes].copy()

    def decode_id_field(self, payload):
        payload = np.ascontiguousarray(payload, dtype=np.uint8)
        n = int(parse_header(payload)["ParticleLen"])
        ids = np.zeros(max(n, 1), np.uint64)
        rc = self.lib.orc_diff_decode_id_field(payload, payload.size, ids)
        if rc:
            raise ValueError("orc_diff_decode_id_field: %d" % rc)
        return ids[:n]

    def encode_uint_field(self, data):
        data = np.ascontiguousarray(data, dtype=np.uint64)
        cap = self.lib.orc_diff_bound((data.size + 2) // 3)
        out = np.zeros(cap, np.uint8)
        nbytes = self.lib.orc_diff_encode_uint_field(data if data.size else np.zeros(1, np.uint64), data.size, out, cap)
        if nbytes < 0:
            raise ValueError("orc_diff_encode_uint_field: %d" % nbytes)
        return out[:nbytes].copy()

    def decode_uint_field(self, payload, n):
        payload = np.ascontiguousarray(payload, dtype=np.uint8)
        out = np.zeros(max(n, 1), np.uint64)
        rc = self.lib.orc_diff_decode_uint_field(payload, payload.size, n, out)
        if rc:
            raise ValueError("orc_diff_decode_uint_field: %d" % rc)
        return out[:n]

    def seg_to_bytes(self, fields, particle_num):
        """fields: [(field_code, particle_len, payload uint8 array)] -> segment image."""
        arr = (SegField * len(fields))()
        keep = []
        for i, (code, n, payload) in enumerate(fields):
            payload = np.ascontiguousarray(payload, dtype=np.uint8)
            keep.append(payload)
            arr[i] = SegField(code, ALGO_DIFF, 0x00000900, n, payload.ctypes.data, payload.size)
        cap = self.lib.orc_seg_bytes_bound(arr, len(fields))
        out = np.zeros(cap, np.uint8)
        nbytes = self.lib.orc_seg_to_bytes(arr, len(fields), particle_num, out, cap)
        if nbytes < 0:
            raise ValueError("orc_seg_to_bytes: %d" % nbytes)
        return out[:nbytes].copy()

    def seg_from_bytes(self, image, max_fields=64):
        """-> ([(field_code, particle_len, payload copy, valid)], particle_num); the image is not modified."""
        buf = np.array(image, dtype=np.uint8, copy=True)
        arr = (SegField * max_fields)()
        valid = (C.c_int32 * max_fields)()
        pn = C.c_int32()
        nf = self.lib.orc_seg_from_bytes(buf, buf.size, arr, max_fields, valid, C.byref(pn))
        if nf < 0:
            raise ValueError("orc_seg_from_bytes: %d" % nf)
        out = []
        for i in range(nf):
            off = arr[i].Data - buf.ctypes.data
            out.append((arr[i].FieldCode, arr[i].ParticleLen, buf[off:off + arr[i].DataLen].copy(), bool(valid[i])))
        return out, pn.value

    def gen_zeldovich(self, ngrid, box, seed, first_id, count, noise_cells=0.0, threads=1):
        """Host twin of the device generator's recipe (not bit-identical): (pos[3*count], vel[3*count])."""
        pos = np.empty(3 * count, np.float32)
        vel = np.empty(3 * count, np.float32)
        rc = self.lib.orc_gen_zeldovich(ngrid, box, seed, first_id, count, noise_cells,
                                        pos.ctypes.data, vel.ctypes.data, threads)
        if rc:
            raise ValueError("orc_gen_zeldovich: %d" % rc)
        return pos, vel

    def bench_encode(self, field_code, data, n, n_fields, delta, width, threads):
        data = np.ascontiguousarray(data, dtype=np.float32).reshape(-1)
        nbytes = C.c_int64()
        secs = self.lib.orc_bench_encode(field_code, data, n, n_fields, delta, width, threads,
                                         C.byref(nbytes))
        return secs, nbytes.value

    def bench_decode(self, payloads, n, threads):
        k = len(payloads)
        ptrs = (C.c_void_p * k)(*[p.ctypes.data for p in payloads])
        lens = (C.c_int64 * k)(*[p.size for p in payloads])
        return self.lib.orc_bench_decode(ptrs, lens, n, k, threads)


def parse_header(payload):
    return np.frombuffer(payload[:64].tobytes(), HEADER_DTYPE)[0]


def parse_table(payload):
    n = int(parse_header(payload)["ParticleLen"])
    nb = (n + BLOCK_LEN - 1) // BLOCK_LEN
    return np.frombuffer(payload[64:64 + 48 * nb].tobytes(), ENTRY_DTYPE)


_cache = {}


def port():
    """The restated oracle (always available)."""
    if "port" not in _cache:
        _cache["port"] = Oracle(build())
    return _cache["port"]


def reference():
    """The same composition on the reference's own primitives, or None."""
    if "ref" not in _cache:
        build()
        path = os.path.join(ORACLE_DIR, "_ref", "liboracle_ref.so")
        _cache["ref"] = Oracle(path) if os.path.exists(path) else None
    return _cache["ref"]


def reference_debug():
    """The reference composition with the reference compiled -D DEBUG (its own
    Makefile's flags), or None. CPU-baseline figure only."""
    if "refdbg" not in _cache:
        build()
        path = os.path.join(ORACLE_DIR, "_ref", "liboracle_ref_debug.so")
        _cache["refdbg"] = Oracle(path) if os.path.exists(path) else None
    return _cache["refdbg"]


def flavours():
    out = [port()]
    if reference() is not None:
        out.append(reference())
    return out
