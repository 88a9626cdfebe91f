"""GPU parity, CSV ingest (App:52-59): dq4ml_csv_parse against the oracle's parser and Python's own float()."""
import numpy as np
import pytest
import torch

from oracle import passes

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", ["abstract", "small", "full"])
@pytest.mark.parametrize("to_device", [True, False])
def test_reference_fixtures(pkg, ctx, lab_csv, golden, name, to_device):
    raw = lab_csv(name)
    g, p = ctx.read_csv(raw, to_device=to_device)
    if to_device:
        g, p = g.cpu().numpy(), p.cpu().numpy()
    doc = golden("dataset_" + name)
    assert g.tolist() == doc["guest"]
    assert p.tobytes() == np.asarray(doc["price"], dtype=np.float64).tobytes()


@pytest.mark.parametrize("sep", [b"\r", b"\n", b"\r\n"])
@pytest.mark.parametrize("n", [1, 7, 1000, 200_003])
def test_synthetic_files_bit_exact(pkg, ctx, sep, n):
    guest, price, _ = passes.synth_generate(n, d=1, p_low=0.05, p_high=0.05)
    recs = [b"%d,%.2f" % (int(a), float(b)) for a, b in zip(guest, price)]
    for trailing in (b"", sep):
        raw = sep.join(recs) + trailing
        og, op = passes.csv_parse(raw)
        g, p = ctx.read_csv(raw)
        assert g.cpu().numpy().tobytes() == og.tobytes()
        assert p.cpu().numpy().tobytes() == op.tobytes()
        assert p.cpu().numpy().tolist() == [float(r.split(b",")[1]) for r in recs]  # == Double.parseDouble


def test_number_forms_and_slow_path(pkg, ctx):
    fields = ["38", "0.1", "-7.25", "+3.5", "1e3", "2.5E-3", "123456789012345", "1234567890123456789", "0.30000000000000004",
              "1.7976931348623157e308", "4.9e-324", "000012.5000", ".5", "5.", "1e23", "9007199254740993", "-0.0", "1e-320"]
    raw = b"\r".join(b"%d,%s" % (i - 3, f.encode()) for i, f in enumerate(fields)) + b"\r\r\n\n"
    g, p = ctx.read_csv(raw)
    assert g.cpu().numpy().tolist() == [i - 3 for i in range(len(fields))]
    exp = np.array([float(f) for f in fields])
    assert p.cpu().numpy().tobytes() == exp.tobytes()


def test_empty_malformed_and_capacity(pkg, ctx):
    import ctypes as C
    g, p = ctx.read_csv(b"")
    assert g.shape[0] == 0 and p.shape[0] == 0
    g, p = ctx.read_csv(b"\r\r\n")
    assert g.shape[0] == 0
    for bad in (b"1;2\r3,4", b"1,2\rx,4\r", b"1,\r", b"1,2,3\r", b"99999999999,2\r"):
        with pytest.raises(pkg.capi.Dq4mlError) as ei:
            ctx.read_csv(bad)
        assert ei.value.status == pkg.capi.ERR_INVALID and "malformed" in str(ei.value)
    # caller's buffers too small
    L = pkg.capi.lib()
    raw = b"1,2\r3,4\r5,6"
    go = torch.zeros(2, dtype=torch.int32, device="cuda")
    po = torch.zeros(2, dtype=torch.float64, device="cuda")
    n = C.c_int64(0)
    buf = C.create_string_buffer(raw, len(raw))
    rc = L.dq4ml_csv_parse(ctx.handle, C.cast(buf, C.c_void_p), len(raw), go.data_ptr(), po.data_ptr(), 2, C.byref(n))
    assert rc == pkg.capi.ERR_INVALID and n.value == 3


def test_device_resident_bytes(pkg, ctx, lab_csv):
    raw = lab_csv("full")
    import ctypes as C
    d = torch.frombuffer(bytearray(raw), dtype=torch.uint8).cuda()
    go = torch.zeros(2000, dtype=torch.int32, device="cuda")
    po = torch.zeros(2000, dtype=torch.float64, device="cuda")
    n = C.c_int64(0)
    rc = pkg.capi.lib().dq4ml_csv_parse(ctx.handle, d.data_ptr(), d.numel(), go.data_ptr(), po.data_ptr(), 2000, C.byref(n))
    assert rc == 0 and n.value == 1040
    og, op = passes.csv_parse(raw)
    assert go[:1040].cpu().numpy().tobytes() == og.tobytes() and po[:1040].cpu().numpy().tobytes() == op.tobytes()
