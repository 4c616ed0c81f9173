"""Host-side C (libminnow_host.so): the registry with the reference's interface
(/root/reference/src/register.h) and the segment flow of src/funcs.c. The CPU
tests only load the library and exercise the registry (no compute calls)."""
import ctypes as C

import numpy as np
import pytest

import minnow_c_b200 as mc


def test_host_library_exports_every_symbol():
    h = mc.host_lib()
    for name in mc.HOST_EXPORTS:
        assert hasattr(h, name), name


def test_registry_knows_algo_diff():
    h = mc.host_lib()
    reg = h.Register_New()
    assert h.Register_Supports(reg, mc.ALGO_DIFF, mc.ALGO_VERSION)
    assert not h.Register_Supports(reg, mc.ALGO_DIFF, 0x00010000)
    assert not h.Register_Supports(reg, 0x5472696D, mc.ALGO_VERSION)  # algo_Test / algo_Trim: a stub upstream
    assert h.Register_Newest(reg, mc.ALGO_DIFF) == mc.ALGO_VERSION
    assert h.Register_Newest(reg, 0x12345678) == 0
    h.Register_Free(reg)


def test_registry_add_is_visible_to_new_registers():
    """Register_Add takes the register by value (src/register.h:28-30); the
    addition must still be seen by the fresh Register_New() every caller in
    src/funcs.c:122-189 makes."""
    h = mc.host_lib()
    reg = h.Register_New()
    funcs = mc.RegisterFuncs()  # null functions: only the lookup is exercised
    h.Register_Add(reg, 0x5A5A5A5A, 0x00010200, funcs)
    h.Register_Add(reg, 0x5A5A5A5A, 0x01010300, funcs)  # stage byte is ignored when comparing
    reg2 = h.Register_New()
    assert h.Register_Supports(reg2, 0x5A5A5A5A, 0x00010200)
    assert h.Register_Newest(reg2, 0x5A5A5A5A) == 0x01010300
    comp = h.Register_GetCompressor(reg2, 0x5A5A5A5A, 0x00010200)
    assert comp.Buffer is None and comp.CFunc is None
    h.Register_FreeCompressor(reg2, 0x5A5A5A5A, 0x00010200, comp)


def test_checksum_golden():
    """util_Checksum({1,2,3}) = 0x60000004 on the compiled reference."""
    h = mc.host_lib()
    data = np.array([1, 2, 3], np.uint8)
    assert h.MnwChecksum(data.ctypes.data, 3) == 0x60000004
    assert h.MnwChecksum(None, 0) == 1


@pytest.mark.gpu
def test_segment_flow_roundtrip_and_damage():
    import _oracle
    from _data import zeldovich_host
    orc = _oracle.port()
    h = mc.host_lib()
    pos, vel = zeldovich_host(16)
    n = pos.shape[1]
    pos, vel = pos.reshape(-1).copy(), vel.reshape(-1).copy()
    pacc = mc.PositionAccuracy(None, 1e-3, 100.0, 0)
    vacc = mc.VelocityAccuracy(None, 0.5, 0, 0, 0.0)
    fields = (mc.Field * 2)()
    fields[0] = mc.Field(mc.FieldHeader(mc.FIELD_POSN, mc.ALGO_DIFF, mc.ALGO_VERSION, n), 1, pos.ctypes.data,
                         C.addressof(pacc))
    fields[1] = mc.Field(mc.FieldHeader(mc.FIELD_VELC, mc.ALGO_DIFF, mc.ALGO_VERSION, n), 1, vel.ctypes.data,
                         C.addressof(vacc))
    h.MnwSeg_SetDitherSeed(99)
    cs = h.MnwSeg_Encode(mc.Seg(fields, 2))
    assert cs.FieldLen == 2
    want = [orc.encode_field(mc.FIELD_POSN, pos, 1e-3, 100.0, seed=99),
            orc.encode_field(mc.FIELD_VELC, vel, 0.5, 0.0, seed=99)]
    for i in range(2):
        cf = cs.Fields[i]
        got = np.ctypeslib.as_array(C.cast(cf.Data, C.POINTER(C.c_uint8)), shape=(cf.DataLen,))
        assert np.array_equal(got, want[i])
        assert cf.Checksum == orc.checksum(want[i])
    # damage the second field: it must come back invalid, the first one intact
    C.cast(cs.Fields[1].Data, C.POINTER(C.c_uint8))[100] ^= 0x40
    s = h.MnwSeg_Decode(cs)
    assert s.Fields[0].Valid == 1 and s.Fields[1].Valid == 0
    got = np.ctypeslib.as_array(C.cast(s.Fields[0].Data, C.POINTER(C.c_float)), shape=(3 * n,))
    ref = orc.decode_field(want[0])
    assert np.array_equal(got.view(np.uint32), ref.view(np.uint32))
    h.MnwSeg_Free(s)
    h.MnwCSeg_Free(cs)
