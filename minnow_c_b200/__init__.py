"""minnow_c_b200 -- Python driver over libminnow_cuda.so (ctypes).

The product is the C-ABI library (include/minnow_cuda.h); this package only
loads it for tests and benchmarks and mirrors the reference's names
(Field / CField / QField, Compress / Decompress, see /root/reference/src/
types.h and funcs.h). torch is used for device memory and streams only.

There is no CPU path: importing works anywhere (so the symbol table can be
checked), but every compute call needs the built library and a CUDA device and
raises loudly otherwise.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(_HERE)
# (MNWCUDA_LIB: kernel experiments load a differently built library of the same ABI)
LIB_PATH = os.environ.get("MNWCUDA_LIB") or os.path.join(_HERE, "libminnow_cuda.so")
HOST_LIB_PATH = os.path.join(_HERE, "libminnow_host.so")
NCCL_LIB_PATH = os.path.join(_HERE, "libminnow_nccl.so")

FIELD_POSN = 0x506F736E
FIELD_VELC = 0x56656C63
FIELD_PTID = 0x50746964
FIELD_UNSI = 0x556E7369
WIDTH_DAMAGED = 0xFFFFFFFF
ALGO_DIFF = 0x44696666
ALGO_VERSION = 0x00000900
BLOCK_LEN = 1024

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-fmad=false",
    "-prec-div=true", "-prec-sqrt=true", "-ftz=false", "-lineinfo", "--extended-lambda",
    "-shared", "-Xcompiler", "-fPIC",
]


def build(verbose=False):
    """Compile libminnow_cuda.so for sm_100a (cross-compiles without a GPU)."""
    src = os.path.join(_HERE, "csrc", "minnow_cuda.cu")
    deps = [os.path.join(_HERE, "csrc", f) for f in os.listdir(os.path.join(_HERE, "csrc"))]
    deps += [os.path.join(ROOT, "include", f) for f in os.listdir(os.path.join(ROOT, "include"))]
    if os.path.exists(LIB_PATH) and all(os.path.getmtime(LIB_PATH) >= os.path.getmtime(d) for d in deps):
        return LIB_PATH
    cmd = ["nvcc"] + NVCC_FLAGS + ["-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(_HERE, "csrc"),
                                   "-o", LIB_PATH, src]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    subprocess.check_call(cmd)
    return LIB_PATH


def build_nccl():
    """Compile libminnow_nccl.so (the C multi-GPU driver; links libminnow_cuda and NCCL)."""
    src = os.path.join(_HERE, "csrc", "minnow_nccl.cu")
    deps = [src, os.path.join(ROOT, "include", "minnow_nccl.h"), os.path.join(ROOT, "include", "minnow_cuda.h")]
    if os.path.exists(NCCL_LIB_PATH) and all(os.path.getmtime(NCCL_LIB_PATH) >= os.path.getmtime(d) for d in deps):
        return NCCL_LIB_PATH
    subprocess.check_call(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo",
                           "-shared", "-Xcompiler", "-fPIC", "-I" + os.path.join(ROOT, "include"), "-o", NCCL_LIB_PATH,
                           src, "-L" + _HERE, "-lminnow_cuda", "-lnccl", "-Xlinker", "-rpath=$ORIGIN"])
    return NCCL_LIB_PATH


# ---- ctypes images of the boundary structs (include/minnow_types.h) ----
class FieldHeader(C.Structure):
    _fields_ = [("FieldCode", C.c_uint32), ("AlgoCode", C.c_uint32), ("AlgoVersion", C.c_uint32),
                ("ParticleLen", C.c_int32)]


class Field(C.Structure):
    _fields_ = [("Hd", FieldHeader), ("Valid", C.c_int64), ("Data", C.c_void_p), ("Acc", C.c_void_p)]


class QField(C.Structure):
    _fields_ = [("Hd", FieldHeader), ("Valid", C.c_int64), ("Data", C.c_void_p), ("Quant", C.c_void_p)]


class CField(C.Structure):
    _fields_ = [("Hd", FieldHeader), ("Data", C.c_void_p), ("DataLen", C.c_int64), ("Checksum", C.c_uint32)]


class PositionAccuracy(C.Structure):
    _fields_ = [("Deltas", C.c_void_p), ("Delta", C.c_float), ("Width", C.c_float), ("Len", C.c_int32)]


class VelocityAccuracy(C.Structure):
    _fields_ = [("Deltas", C.c_void_p), ("Delta", C.c_float), ("Len", C.c_int32),
                ("SymLog10Scaled", C.c_int32), ("SymLog10Threshold", C.c_float)]


class PositionQuantization(C.Structure):
    _fields_ = [("Depths", C.c_void_p), ("Len", C.c_int32), ("Width", C.c_float), ("X0", C.c_float * 3),
                ("X1", C.c_float * 3), ("Depth", C.c_uint8)]


class VelocityQuantization(C.Structure):
    _fields_ = [("Depths", C.c_void_p), ("Len", C.c_int32), ("SymLog10Scaled", C.c_int32),
                ("X0", C.c_float * 3), ("X1", C.c_float * 3), ("SymLog10Threshold", C.c_float),
                ("Depth", C.c_uint8)]


class IDAccuracy(C.Structure):
    _fields_ = [("Width", C.c_uint64)]


class IDQuantization(C.Structure):
    _fields_ = [("Width", C.c_uint64), ("X0", C.c_uint64 * 3), ("X1", C.c_uint64 * 3)]


class IntQuantization(C.Structure):
    _fields_ = [("X0", C.c_uint64), ("X1", C.c_uint64)]


class BlockHeader(C.Structure):
    _fields_ = [("Length", C.c_int32), ("Checksum", C.c_uint32)]


class FieldDesc(C.Structure):
    _fields_ = [("Offset", C.c_int64), ("ParticleLen", C.c_int32), ("FieldCode", C.c_uint32),
                ("Delta", C.c_float), ("Width", C.c_float), ("DitherSeed", C.c_uint64)]


EXPORTS = {
    "mnwcuda_FieldFixedBytes": (C.c_int64, [C.c_int32]),
    "mnwcuda_FieldBound": (C.c_int64, [C.c_int32]),
    "mnwcuda_Create": (C.c_void_p, [C.c_int]),
    "mnwcuda_Destroy": (None, [C.c_void_p]),
    "mnwcuda_EncodeFieldsDev": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(FieldDesc), C.c_int64, C.c_void_p,
                                          C.c_int64, C.c_void_p, C.c_void_p]),
    "mnwcuda_DecodeFieldsDev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(FieldDesc), C.c_int64,
                                          C.c_void_p, C.c_void_p]),
    "mnwcuda_Status": (C.c_int, [C.c_void_p, C.c_void_p]),
    "mnwcuda_LaunchCount": (C.c_int64, [C.c_void_p]),
    "mnwcuda_EnableKernelTiming": (None, [C.c_void_p, C.c_int]),
    "mnwcuda_KernelTimes": (C.c_int, [C.c_void_p, C.POINTER(C.c_float)]),
    "mnwcuda_EncodeField": (CField, [Field, C.c_void_p]),
    "mnwcuda_DecodeField": (Field, [CField, C.c_void_p]),
    "mnwcuda_EncodeFieldsHost": (C.c_int64, [C.c_void_p, C.c_void_p, C.POINTER(FieldDesc), C.c_int64, C.c_void_p,
                                             C.c_int64, C.c_void_p]),
    "mnwcuda_DecodeFieldsHost": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(FieldDesc), C.c_int64,
                                           C.c_void_p]),
    "MnwCudaCompress_v0_9": (CField, [QField, C.c_void_p]),
    "MnwCudaDecompress_v0_9": (QField, [CField, C.c_void_p]),
    "MnwCudaCAlloc_v0_9": (C.c_void_p, []),
    "MnwCudaCFree_v0_9": (None, [C.c_void_p]),
    "MnwCudaDAlloc_v0_9": (C.c_void_p, []),
    "MnwCudaDFree_v0_9": (None, [C.c_void_p]),
    "mnwcuda_SetDitherSeed": (None, [C.c_void_p, C.c_uint64]),
    "mnwcuda_GenerateZeldovich": (C.c_int, [C.c_void_p, C.c_int32, C.c_float, C.c_uint64, C.c_int64, C.c_int64,
                                            C.c_void_p, C.c_void_p, C.c_void_p]),
    "mnwcuda_Acquire": (C.c_void_p, [C.c_int]),
    "mnwcuda_Release": (None, [C.c_void_p]),
    "mnwcuda_QuantizeField": (QField, [Field, C.c_void_p]),
    "mnwcuda_DequantizeField": (Field, [QField, C.c_void_p]),
    "mnwcuda_ForgetQuant": (None, [C.c_void_p]),
    "mnwcuda_BlockCount": (C.c_int64, [C.POINTER(FieldDesc), C.c_int64]),
    "mnwcuda_BlockHeadersDev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(FieldDesc), C.c_int64,
                                          C.c_void_p, C.c_void_p]),
    "mnwcuda_VerifyBlocksDev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(FieldDesc), C.c_int64,
                                          C.c_void_p, C.c_void_p, C.c_void_p]),
    "mnwcuda_ChecksumsHost": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int64,
                                        C.c_void_p]),
    "mnwcuda_SegmentImageBound": (C.c_int64, [C.POINTER(CField), C.c_int32]),
    "mnwcuda_SegmentToImage": (C.c_int64, [C.c_void_p, C.POINTER(CField), C.c_int32, C.c_int32, C.c_void_p, C.c_int64]),
    "mnwcuda_SegmentFromImage": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_int64, C.POINTER(CField), C.c_int32,
                                             C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "mnwcuda_GenerateZeldovichRough": (C.c_int, [C.c_void_p, C.c_int32, C.c_float, C.c_uint64, C.c_int64, C.c_int64,
                                                 C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]),
    "mnwcuda_DecodeFieldNoAlloc": (C.c_int, [C.c_int, C.c_void_p, C.c_int64, C.c_int32, C.c_uint32, C.c_void_p,
                                             C.c_void_p, C.c_void_p]),
}
NOALLOC_MAX_PARTICLES = 1 << 22  # MNWCUDA_NOALLOC_MAX_PARTICLES

class _Seq(C.Structure):
    _fields_ = [("Data", C.c_void_p), ("Len", C.c_int32), ("Cap", C.c_int32)]


class Register(C.Structure):  # include/minnow_register.h (src/register.h:9-13)
    _fields_ = [("Algo", _Seq), ("Version", _Seq), ("Funcs", _Seq)]


class RegisterFuncs(C.Structure):  # src/register.h:15-22
    _fields_ = [("CFunc", C.c_void_p), ("CAlloc", C.c_void_p), ("CFree", C.c_void_p),
                ("DFunc", C.c_void_p), ("DAlloc", C.c_void_p), ("DFree", C.c_void_p)]


class Compressor(C.Structure):
    _fields_ = [("Buffer", C.c_void_p), ("CFunc", C.c_void_p)]


class Decompressor(C.Structure):
    _fields_ = [("Buffer", C.c_void_p), ("DFunc", C.c_void_p)]


class Seg(C.Structure):
    _fields_ = [("Fields", C.POINTER(Field)), ("FieldLen", C.c_int32)]


class CSeg(C.Structure):
    _fields_ = [("Fields", C.POINTER(CField)), ("FieldLen", C.c_int32)]


class QSeg(C.Structure):
    _fields_ = [("Fields", C.POINTER(QField)), ("FieldLen", C.c_int32)]


class Bytes(C.Structure):  # MnwBytes
    _fields_ = [("Data", C.c_void_p), ("Len", C.c_int64)]


HOST_EXPORTS = {
    "Register_New": (Register, []),
    "Register_Free": (None, [Register]),
    "Register_Add": (None, [Register, C.c_uint32, C.c_uint32, RegisterFuncs]),
    "Register_GetCompressor": (Compressor, [Register, C.c_uint32, C.c_uint32]),
    "Register_GetDecompressor": (Decompressor, [Register, C.c_uint32, C.c_uint32]),
    "Register_FreeCompressor": (None, [Register, C.c_uint32, C.c_uint32, Compressor]),
    "Register_FreeDecompressor": (None, [Register, C.c_uint32, C.c_uint32, Decompressor]),
    "Register_Supports": (C.c_bool, [Register, C.c_uint32, C.c_uint32]),
    "Register_Newest": (C.c_uint32, [Register, C.c_uint32]),
    "MnwSeg_Encode": (CSeg, [Seg]),
    "MnwSeg_Decode": (Seg, [CSeg]),
    "MnwCSeg_Free": (None, [CSeg]),
    "MnwSeg_Free": (None, [Seg]),
    "MnwChecksum": (C.c_uint32, [C.c_void_p, C.c_int64]),
    "MnwSeg_SetDitherSeed": (None, [C.c_uint64]),
    "MnwSeg_ToBytes": (Bytes, [CSeg, C.c_int32]),
    "MnwSeg_FromBytes": (CSeg, [Bytes, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "quant_QField": (QField, [Field]),
    "quant_Field": (Field, [QField]),
    "quant_FreeQField": (None, [QField]),
    "quant_FreeField": (None, [Field]),
    "MnwQuant_SetDitherSeed": (None, [C.c_uint64]),
}

NCCL_EXPORTS = {
    "mnwnccl_UniqueId": (C.c_int, [C.c_void_p]),
    "mnwnccl_Init": (C.c_void_p, [C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "mnwnccl_Finalize": (None, [C.c_void_p]),
    "mnwnccl_Rank": (C.c_int, [C.c_void_p]),
    "mnwnccl_World": (C.c_int, [C.c_void_p]),
    "mnwnccl_EncodeSharded": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(FieldDesc), C.c_int64, C.c_void_p, C.c_int64,
                                        C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "mnwnccl_DecodeSharded": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(FieldDesc), C.c_int64, C.c_void_p,
                                        C.c_void_p]),
    "mnwnccl_GatherStream": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "mnwnccl_ScatterStream": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
}

_lib = None
_host = None
_nccl = None
_libc = C.CDLL(None)
_libc.free.argtypes = [C.c_void_p]


def lib():
    """The loaded libminnow_cuda.so. Raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("libminnow_cuda.so is not built (run __graft_entry__.build()); "
                               "there is no CPU fallback")
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in EXPORTS.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def host_lib():
    """The loaded libminnow_host.so (registry + segment flow, host-side C)."""
    global _host
    if _host is None:
        lib()  # libminnow_cuda.so first: the host library links against it
        if not os.path.exists(HOST_LIB_PATH):
            raise RuntimeError("libminnow_host.so is not built (run __graft_entry__.build())")
        handle = C.CDLL(HOST_LIB_PATH)
        for name, (res, args) in HOST_EXPORTS.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _host = handle
    return _host


def nccl_lib():
    """The loaded libminnow_nccl.so (C multi-GPU driver). torch is imported first so that
    its bundled libnccl.so.2 is the one both share."""
    global _nccl
    if _nccl is None:
        import torch  # noqa: F401
        lib()
        if not os.path.exists(NCCL_LIB_PATH):
            raise RuntimeError("libminnow_nccl.so is not built (run __graft_entry__.build())")
        handle = C.CDLL(NCCL_LIB_PATH)
        for name, (res, args) in NCCL_EXPORTS.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _nccl = handle
    return _nccl


class Sharded:
    """One rank of the C multi-GPU driver (include/minnow_nccl.h) over a Codec."""

    def __init__(self, codec, rank, world, unique_id):
        self.codec, self.rank, self.world = codec, rank, world
        buf = (C.c_uint8 * 128).from_buffer_copy(bytes(unique_id))
        self.h = nccl_lib().mnwnccl_Init(codec.ctx, rank, world, buf)
        if not self.h:
            raise RuntimeError("mnwnccl_Init failed")

    @staticmethod
    def unique_id():
        buf = (C.c_uint8 * 128)()
        if nccl_lib().mnwnccl_UniqueId(buf):
            raise RuntimeError("mnwnccl_UniqueId failed")
        return bytes(buf)

    def close(self):
        if self.h:
            nccl_lib().mnwnccl_Finalize(self.h)
            self.h = None

    def encode(self, d_in, descs, n_fields, d_out, d_offsets, d_sizes, d_base, stream=None):
        rc = nccl_lib().mnwnccl_EncodeSharded(self.h, d_in.data_ptr(), descs, n_fields, d_out.data_ptr(),
                                              d_out.numel() * d_out.element_size(), d_offsets.data_ptr(),
                                              d_sizes.data_ptr(), d_base.data_ptr(), _stream_ptr(stream))
        if rc:
            raise ValueError("mnwnccl_EncodeSharded: %d" % rc)

    def decode(self, d_in, d_offsets, descs, n_fields, d_out, stream=None):
        rc = nccl_lib().mnwnccl_DecodeSharded(self.h, d_in.data_ptr(), d_offsets.data_ptr(), descs, n_fields,
                                              d_out.data_ptr(), _stream_ptr(stream))
        if rc:
            raise ValueError("mnwnccl_DecodeSharded: %d" % rc)

    def gather(self, d_local, sizes, root, d_global=None, stream=None):
        arr = (C.c_uint64 * self.world)(*[int(v) for v in sizes])
        rc = nccl_lib().mnwnccl_GatherStream(self.h, d_local.data_ptr(), arr, root,
                                             d_global.data_ptr() if d_global is not None else None, _stream_ptr(stream))
        if rc:
            raise ValueError("mnwnccl_GatherStream: %d" % rc)

    def scatter(self, d_global, sizes, root, d_local, stream=None):
        arr = (C.c_uint64 * self.world)(*[int(v) for v in sizes])
        rc = nccl_lib().mnwnccl_ScatterStream(self.h, d_global.data_ptr() if d_global is not None else None, arr, root,
                                              d_local.data_ptr(), _stream_ptr(stream))
        if rc:
            raise ValueError("mnwnccl_ScatterStream: %d" % rc)


def field_bound(n):
    return lib().mnwcuda_FieldBound(n)


def field_fixed_bytes(n):
    return lib().mnwcuda_FieldFixedBytes(n)


def make_descs(fields):
    """fields: iterable of (offset, n, field_code, delta, width, seed)."""
    fields = list(fields)
    arr = (FieldDesc * max(len(fields), 1))()
    for i, (off, n, code, delta, width, seed) in enumerate(fields):
        arr[i] = FieldDesc(off, n, code, delta, width, seed)
    return arr


def _stream_ptr(stream):
    if stream is None:
        import torch
        return C.c_void_p(torch.cuda.current_stream().cuda_stream)
    return C.c_void_p(stream)


class Codec:
    """One MnwCudaCtx: device scratch + the entry points, numpy/torch friendly."""

    def __init__(self, device=None):
        import torch
        if not torch.cuda.is_available():
            raise RuntimeError("minnow_c_b200 needs a CUDA device; there is no CPU fallback")
        if device is None:
            device = torch.cuda.current_device()
        self.device = int(device)
        self.ctx = lib().mnwcuda_Create(self.device)
        if not self.ctx:
            raise RuntimeError("mnwcuda_Create failed")

    def close(self):
        if self.ctx:
            lib().mnwcuda_Destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def launches(self):
        return lib().mnwcuda_LaunchCount(self.ctx)

    # mnwcuda_KernelTimes slots; None = not used by this version of the library
    KERNEL_NAMES = ("k_encode", "k_finalize+k_tables", "k_block_headers", None,
                    "k_parse_headers", "k_stream_states", "k_unpack_dequant")

    def enable_kernel_timing(self, on=True):
        lib().mnwcuda_EnableKernelTiming(self.ctx, 1 if on else 0)

    def kernel_times(self):
        """ms per kernel of the most recent encode_dev / decode_dev (waits for them)."""
        ms = (C.c_float * 7)()
        if lib().mnwcuda_KernelTimes(self.ctx, ms):
            raise RuntimeError("kernel timing is not enabled")
        return {k: float(v) for k, v in zip(self.KERNEL_NAMES, ms) if k is not None}

    # ---- device-resident batch path ----
    def encode_dev(self, d_in, descs, n_fields, d_out, d_offsets, stream=None):
        rc = lib().mnwcuda_EncodeFieldsDev(self.ctx, d_in.data_ptr(), descs, n_fields, d_out.data_ptr(),
                                           d_out.numel() * d_out.element_size(), d_offsets.data_ptr(),
                                           _stream_ptr(stream))
        if rc:
            raise ValueError("mnwcuda_EncodeFieldsDev: %d" % rc)

    def decode_dev(self, d_in, d_offsets, descs, n_fields, d_out, stream=None):
        rc = lib().mnwcuda_DecodeFieldsDev(self.ctx, d_in.data_ptr(), d_offsets.data_ptr(), descs, n_fields,
                                           d_out.data_ptr(), _stream_ptr(stream))
        if rc:
            raise ValueError("mnwcuda_DecodeFieldsDev: %d" % rc)

    def status(self, stream=None):
        return lib().mnwcuda_Status(self.ctx, _stream_ptr(stream))

    def generate(self, ngrid, box, seed, first_id, count, d_pos=None, d_vel=None, stream=None, noise_cells=0.0):
        rc = lib().mnwcuda_GenerateZeldovichRough(self.ctx, ngrid, box, seed, first_id, count, noise_cells,
                                                  d_pos.data_ptr() if d_pos is not None else None,
                                                  d_vel.data_ptr() if d_vel is not None else None,
                                                  _stream_ptr(stream))
        if rc:
            raise ValueError("mnwcuda_GenerateZeldovich: %d" % rc)

    # ---- host-buffer batch path ----
    def encode_host(self, h_in, descs, n_fields, h_out, h_offsets):
        total = lib().mnwcuda_EncodeFieldsHost(self.ctx, _host_ptr(h_in), descs, n_fields, _host_ptr(h_out),
                                               _host_nbytes(h_out), _host_ptr(h_offsets))
        if total < 0:
            raise ValueError("mnwcuda_EncodeFieldsHost: %d" % total)
        return total

    def decode_host(self, h_in, h_offsets, descs, n_fields, h_out):
        rc = lib().mnwcuda_DecodeFieldsHost(self.ctx, _host_ptr(h_in), _host_ptr(h_offsets), descs, n_fields,
                                            _host_ptr(h_out))
        if rc:
            raise ValueError("mnwcuda_DecodeFieldsHost: %d" % rc)

    # ---- reference-shaped single-field calls (host numpy in/out) ----
    def set_dither_seed(self, seed):
        lib().mnwcuda_SetDitherSeed(self.ctx, seed)

    def encode_field(self, field_code, data, delta, width=0.0):
        """mnwcuda_EncodeField: float field (3N planar) -> payload bytes (numpy uint8)."""
        data = np.ascontiguousarray(data, dtype=np.float32).reshape(-1)
        n = data.size // 3
        if field_code == FIELD_POSN:
            acc = PositionAccuracy(None, delta, width, 0)
        else:
            acc = VelocityAccuracy(None, delta, 0, 0, 0.0)
        f = Field(FieldHeader(field_code, ALGO_DIFF, ALGO_VERSION, n), 1, data.ctypes.data, C.addressof(acc))
        cf = lib().mnwcuda_EncodeField(f, self.ctx)
        out = np.ctypeslib.as_array(C.cast(cf.Data, C.POINTER(C.c_uint8)), shape=(cf.DataLen,)).copy()
        _libc.free(cf.Data)
        return out

    def decode_field(self, field_code, n, payload):
        """mnwcuda_DecodeField: payload -> (floats[3N], reported Delta)."""
        payload = np.ascontiguousarray(payload, dtype=np.uint8)
        cf = CField(FieldHeader(field_code, ALGO_DIFF, ALGO_VERSION, n), payload.ctypes.data, payload.size, 0)
        f = lib().mnwcuda_DecodeField(cf, self.ctx)
        if n:
            out = np.ctypeslib.as_array(C.cast(f.Data, C.POINTER(C.c_float)), shape=(3 * n,)).copy()
        else:
            out = np.zeros(0, np.float32)
        if field_code == FIELD_POSN:
            delta = C.cast(f.Acc, C.POINTER(PositionAccuracy)).contents.Delta
        else:
            delta = C.cast(f.Acc, C.POINTER(VelocityAccuracy)).contents.Delta
        _libc.free(f.Data)
        _libc.free(f.Acc)
        return out, delta

    # ---- quantiser alone and integer fields ----
    def _field(self, field_code, data, delta=0.0, width=0.0):
        """(Field, keep-alive objects) for a host array"""
        if field_code in (FIELD_POSN, FIELD_VELC):
            data = np.ascontiguousarray(data, dtype=np.float32).reshape(-1)
            n = data.size // 3
            acc = PositionAccuracy(None, delta, width, 0) if field_code == FIELD_POSN else \
                VelocityAccuracy(None, delta, 0, 0, 0.0)
        else:
            data = np.ascontiguousarray(data, dtype=np.uint64).reshape(-1)
            n = data.size
            acc = IDAccuracy(int(width)) if field_code == FIELD_PTID else C.c_uint8(0)
        f = Field(FieldHeader(field_code, ALGO_DIFF, ALGO_VERSION, n), 1, data.ctypes.data, C.addressof(acc))
        return f, (data, acc)

    def quantize_field(self, field_code, data, delta=0.0, width=0.0):
        """mnwcuda_QuantizeField -> (q uint64 array, quant dict)"""
        f, keep = self._field(field_code, data, delta, width)
        qf = lib().mnwcuda_QuantizeField(f, self.ctx)
        n = f.Hd.ParticleLen
        count = n if field_code == FIELD_UNSI else 3 * n
        q = np.ctypeslib.as_array(C.cast(qf.Data, C.POINTER(C.c_uint64)), shape=(count,)).copy() if count else \
            np.zeros(0, np.uint64)
        info = _quant_info(field_code, qf.Quant)
        _libc.free(qf.Data)
        _libc.free(qf.Quant)
        return q, info

    def dequantize_field(self, field_code, n, q64, info):
        """mnwcuda_DequantizeField: (q, quant dict) -> field data (float32[3n] or uint64[n])"""
        q64 = np.ascontiguousarray(q64, dtype=np.uint64)
        quant = _quant_struct(field_code, info)
        qf = QField(FieldHeader(field_code, ALGO_DIFF, ALGO_VERSION, n), 1, q64.ctypes.data, C.addressof(quant))
        f = lib().mnwcuda_DequantizeField(qf, self.ctx)
        if field_code in (FIELD_POSN, FIELD_VELC):
            out = np.ctypeslib.as_array(C.cast(f.Data, C.POINTER(C.c_float)), shape=(3 * n,)).copy() if n else \
                np.zeros(0, np.float32)
        else:
            out = np.ctypeslib.as_array(C.cast(f.Data, C.POINTER(C.c_uint64)), shape=(n,)).copy() if n else \
                np.zeros(0, np.uint64)
        _libc.free(f.Data)
        if f.Acc:
            _libc.free(f.Acc)
        return out

    def encode_ints(self, field_code, data, width=0):
        """mnwcuda_EncodeField for Ptid (width = grid cells per side) / Unsi -> payload bytes"""
        f, keep = self._field(field_code, data, 0.0, width)
        cf = lib().mnwcuda_EncodeField(f, self.ctx)
        out = np.ctypeslib.as_array(C.cast(cf.Data, C.POINTER(C.c_uint8)), shape=(cf.DataLen,)).copy()
        _libc.free(cf.Data)
        return out

    def decode_ints(self, field_code, n, payload):
        payload = np.ascontiguousarray(payload, dtype=np.uint8)
        cf = CField(FieldHeader(field_code, ALGO_DIFF, ALGO_VERSION, n), payload.ctypes.data, payload.size, 0)
        f = lib().mnwcuda_DecodeField(cf, self.ctx)
        out = np.ctypeslib.as_array(C.cast(f.Data, C.POINTER(C.c_uint64)), shape=(n,)).copy() if n else \
            np.zeros(0, np.uint64)
        _libc.free(f.Data)
        if f.Acc:
            _libc.free(f.Acc)
        return out

    def block_headers_dev(self, d_stream, d_offsets, descs, n_fields, stream=None):
        """mnwcuda_BlockHeadersDev -> torch tensor of int32 pairs [nBlocks, 2] on the device"""
        import torch
        nblk = lib().mnwcuda_BlockCount(descs, n_fields)
        d_hdr = torch.empty((nblk, 2), dtype=torch.int32, device=d_stream.device)
        rc = lib().mnwcuda_BlockHeadersDev(self.ctx, d_stream.data_ptr(), d_offsets.data_ptr(), descs, n_fields,
                                           d_hdr.data_ptr(), _stream_ptr(stream))
        if rc:
            raise ValueError("mnwcuda_BlockHeadersDev: %d" % rc)
        return d_hdr

    def compress(self, field_code, q64, x0, x1, depth, width=0.0):
        """MnwCudaCompress_v0_9: the literal CFunc(QField) boundary."""
        q64 = np.ascontiguousarray(q64, dtype=np.uint64)
        n = q64.size // 3
        if field_code == FIELD_POSN:
            quant = PositionQuantization(None, 0, width, (C.c_float * 3)(*x0), (C.c_float * 3)(*x1), depth)
        else:
            quant = VelocityQuantization(None, 0, 0, (C.c_float * 3)(*x0), (C.c_float * 3)(*x1), 0.0, depth)
        qf = QField(FieldHeader(field_code, ALGO_DIFF, ALGO_VERSION, n), 1, q64.ctypes.data, C.addressof(quant))
        cf = lib().MnwCudaCompress_v0_9(qf, self.ctx)
        out = np.ctypeslib.as_array(C.cast(cf.Data, C.POINTER(C.c_uint8)), shape=(cf.DataLen,)).copy()
        _libc.free(cf.Data)
        return out

    def decompress(self, field_code, n, payload):
        """MnwCudaDecompress_v0_9: the literal DFunc(CField) boundary -> (q64, quant dict)."""
        payload = np.ascontiguousarray(payload, dtype=np.uint8)
        cf = CField(FieldHeader(field_code, ALGO_DIFF, ALGO_VERSION, n), payload.ctypes.data, payload.size, 0)
        qf = lib().MnwCudaDecompress_v0_9(cf, self.ctx)
        q = np.ctypeslib.as_array(C.cast(qf.Data, C.POINTER(C.c_uint64)), shape=(3 * n,)).copy() if n else \
            np.zeros(0, np.uint64)
        cls = PositionQuantization if field_code == FIELD_POSN else VelocityQuantization
        qt = C.cast(qf.Quant, C.POINTER(cls)).contents
        info = {"X0": list(qt.X0), "X1": list(qt.X1), "Depth": int(qt.Depth)}
        if field_code == FIELD_POSN:
            info["Width"] = float(qt.Width)
        _libc.free(qf.Data)
        _libc.free(qf.Quant)
        return q, info


def _quant_info(field_code, ptr):
    if field_code == FIELD_POSN:
        q = C.cast(ptr, C.POINTER(PositionQuantization)).contents
        return {"X0": list(q.X0), "X1": list(q.X1), "Depth": int(q.Depth), "Width": float(q.Width)}
    if field_code == FIELD_VELC:
        q = C.cast(ptr, C.POINTER(VelocityQuantization)).contents
        return {"X0": list(q.X0), "X1": list(q.X1), "Depth": int(q.Depth)}
    if field_code == FIELD_PTID:
        q = C.cast(ptr, C.POINTER(IDQuantization)).contents
        return {"X0": list(q.X0), "X1": list(q.X1), "Width": int(q.Width)}
    q = C.cast(ptr, C.POINTER(IntQuantization)).contents
    return {"X0": int(q.X0), "X1": int(q.X1)}


def _quant_struct(field_code, info):
    if field_code == FIELD_POSN:
        return PositionQuantization(None, 0, info["Width"], (C.c_float * 3)(*info["X0"]), (C.c_float * 3)(*info["X1"]),
                                    info["Depth"])
    if field_code == FIELD_VELC:
        return VelocityQuantization(None, 0, 0, (C.c_float * 3)(*info["X0"]), (C.c_float * 3)(*info["X1"]), 0.0,
                                    info["Depth"])
    if field_code == FIELD_PTID:
        return IDQuantization(info["Width"], (C.c_uint64 * 3)(*info["X0"]), (C.c_uint64 * 3)(*info["X1"]))
    return IntQuantization(info["X0"], info["X1"])


def _host_ptr(a):
    if hasattr(a, "data_ptr"):
        return a.data_ptr()
    return a.ctypes.data


def _host_nbytes(a):
    if hasattr(a, "data_ptr"):
        return a.numel() * a.element_size()
    return a.nbytes
