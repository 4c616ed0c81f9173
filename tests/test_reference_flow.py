"""GPU: the reference's OWN segment flow -- /root/reference/src/funcs.c compiled
from where it lies (oracle/Makefile, only its example main() renamed) against
libminnow_host.so: Quantize -> LoadCompressors -> Compress -> Decompress ->
UndoQuantize (src/funcs.c:13-76, 122-189) with AlgoCode = algo_Diff, every
function the reference's, registry / quantiser / plugin functions the product's.
The library is built in the container (the sources do not travel) and rides to
the GPU box in oracle/_ref/."""
import ctypes as C
import os

import numpy as np
import pytest

import _oracle
from _data import zeldovich_host

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
import minnow_c_b200 as mc  # noqa: E402

FUNCS = os.path.join(mc.ROOT, "oracle", "_ref", "libminnow_funcs.so")


@pytest.fixture(scope="module")
def funcs():
    if not os.path.exists(FUNCS):
        pytest.skip("oracle/_ref/libminnow_funcs.so was not built (needs /root/reference at build time)")
    mc.host_lib()  # dependencies first
    lib = C.CDLL(FUNCS)
    lib.Quantize.restype, lib.Quantize.argtypes = mc.QSeg, [mc.Seg]
    lib.UndoQuantize.restype, lib.UndoQuantize.argtypes = mc.Seg, [mc.QSeg]
    lib.Compress.restype, lib.Compress.argtypes = mc.CSeg, [mc.QSeg, C.c_void_p]
    lib.Decompress.restype, lib.Decompress.argtypes = mc.QSeg, [mc.CSeg, C.c_void_p]
    lib.LoadCompressors.restype, lib.LoadCompressors.argtypes = C.c_void_p, [mc.Seg]
    lib.LoadDecompressors.restype, lib.LoadDecompressors.argtypes = C.c_void_p, [mc.CSeg]
    lib.FreeCompressors.restype, lib.FreeCompressors.argtypes = None, [mc.Seg, C.c_void_p]
    lib.FreeDecompressors.restype, lib.FreeDecompressors.argtypes = None, [mc.CSeg, C.c_void_p]
    lib.QSeg_Free.restype, lib.QSeg_Free.argtypes = None, [mc.QSeg]
    lib.CSeg_Free.restype, lib.CSeg_Free.argtypes = None, [mc.CSeg]
    lib.Seg_Free.restype, lib.Seg_Free.argtypes = None, [mc.Seg]
    return lib


@pytest.mark.parametrize("orc", _oracle.flavours(), ids=[o.flavour for o in _oracle.flavours()])
def test_reference_flow_runs_on_the_cuda_plugin(funcs, orc):
    pos, vel = zeldovich_host(32)
    n = pos.shape[1]
    pos, vel = pos.reshape(-1).copy(), vel.reshape(-1).copy()
    ids = ((np.arange(n, dtype=np.uint64) + np.uint64(3077)) % np.uint64(n))  # all below Width^3, wrapping in z
    pacc = mc.PositionAccuracy(None, 1e-3, 100.0, 0)
    vacc = mc.VelocityAccuracy(None, 1.0, 0, 0, 0.0)
    iacc = mc.IDAccuracy(32)
    fields = (mc.Field * 3)()
    fields[0] = mc.Field(mc.FieldHeader(mc.FIELD_POSN, mc.ALGO_DIFF, mc.ALGO_VERSION, n), 1, pos.ctypes.data, C.addressof(pacc))
    fields[1] = mc.Field(mc.FieldHeader(mc.FIELD_VELC, mc.ALGO_DIFF, mc.ALGO_VERSION, n), 1, vel.ctypes.data, C.addressof(vacc))
    fields[2] = mc.Field(mc.FieldHeader(mc.FIELD_PTID, mc.ALGO_DIFF, mc.ALGO_VERSION, n), 1, ids.ctypes.data, C.addressof(iacc))
    seg = mc.Seg(fields, 3)

    qs = funcs.Quantize(seg)                     # src/funcs.c:13-23 -> quant_QField
    assert qs.FieldLen == 3
    hd, q = orc.quantize_field(mc.FIELD_POSN, pos, 1e-3, 100.0)
    got_q = np.ctypeslib.as_array(C.cast(qs.Fields[0].Data, C.POINTER(C.c_uint64)), shape=(3 * n,))
    assert np.array_equal(got_q, q.astype(np.uint64))
    assert qs.Fields[0].Quant  # the Quantization is attached (the reference forgets to)

    comps = funcs.LoadCompressors(seg)           # src/funcs.c:143-162 -> Register_GetCompressor -> CAlloc
    cs = funcs.Compress(qs, comps)               # src/funcs.c:62-76 -> CFunc + util_Checksum
    want = [orc.encode_field(mc.FIELD_POSN, pos, 1e-3, 100.0, seed=0),
            orc.encode_field(mc.FIELD_VELC, vel, 1.0, 0.0, seed=0),
            orc.encode_id_field(ids, 32)]
    for i in range(3):
        cf = cs.Fields[i]
        payload = np.ctypeslib.as_array(C.cast(cf.Data, C.POINTER(C.c_uint8)), shape=(cf.DataLen,))
        assert np.array_equal(payload, want[i]), "field %d" % i
        assert cf.Checksum == orc.checksum(want[i])
    funcs.QSeg_Free(qs)                          # src/funcs.c:198-203 -> quant_FreeQField

    # damage the velocity field: Decompress must skip it (src/funcs.c:49-56)
    C.cast(cs.Fields[1].Data, C.POINTER(C.c_uint8))[200] ^= 0x04
    decomps = funcs.LoadDecompressors(cs)        # src/funcs.c:122-141
    qs2 = funcs.Decompress(cs, decomps)          # -> DFunc
    assert qs2.Fields[0].Valid == 1 and qs2.Fields[1].Valid == 0 and qs2.Fields[2].Valid == 1
    back = funcs.UndoQuantize(qs2)               # src/funcs.c:25-38 -> quant_Field
    f0 = np.ctypeslib.as_array(C.cast(back.Fields[0].Data, C.POINTER(C.c_float)), shape=(3 * n,))
    assert np.array_equal(f0.view(np.uint32), orc.decode_field(want[0]).view(np.uint32))
    f2 = np.ctypeslib.as_array(C.cast(back.Fields[2].Data, C.POINTER(C.c_uint64)), shape=(n,))
    assert np.array_equal(f2, ids)
    assert not back.Fields[1].Data and back.Fields[1].Valid == 0

    funcs.FreeCompressors(seg, comps)
    funcs.FreeDecompressors(cs, decomps)
    funcs.CSeg_Free(cs)
