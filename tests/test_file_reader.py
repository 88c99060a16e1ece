"""The fast matrix-file reader (csrc/fileio.hpp) against the reference's own
EigSol::readMatrixFromFile (src/reader/file_matrix_reader.hpp:170-200), compiled from the unmodified reference
into oracle/_ref: same arrays bit for bit on the reference's data files and on generated dense / sparse,
real / complex files, and the reference's error messages on malformed input.  Where oracle/_ref is not
available (the GPU box never needs it: nothing here is gpu-marked) the format is checked against numpy."""
import os

import numpy as np
import pytest

import pcsc_eigenvalue_solver_project_b200 as es
from pcsc_eigenvalue_solver_project_b200 import generators as gen
from _oracle import Reference, have_reference

REF_DATA = "/root/reference/data"


@pytest.fixture(scope="module")
def ref():
    if not have_reference():
        pytest.skip("oracle/_ref not built")
    r = Reference()
    if not hasattr(r.L, "ref_read_matrix_file"):
        pytest.skip("oracle/_ref predates the reader shim")
    return r


def _fmt(v, cplx):
    v = complex(v)
    return f"{v.real!r} {v.imag!r}" if cplx else repr(v.real)


def write_sparse(path, rows, cols, entries, cplx, shuffle_ws=False):
    with open(path, "w") as f:
        f.write(f"sparse\n{rows} {cols}\n{len(entries)}\n")
        for k, (r, c, v) in enumerate(entries):
            sep = "\t \n"[k % 3] if shuffle_ws else " "
            f.write(f"{r}{sep}{c} {_fmt(v, cplx)}\n")


def write_dense(path, a, cplx):
    with open(path, "w") as f:
        f.write(f"dense {a.shape[0]} {a.shape[1]}\n")
        for row in a:
            f.write("  ".join(_fmt(v, cplx) for v in row) + "\n")


def same_as_reference(ref, path, dtype):
    m = es.readMatrixFromFile(path, dtype)
    isd, rows, cols, outer, inner, vals, _ = ref.read_matrix_file(path, dtype)
    assert m.isDense() == isd and m._shape == (rows, cols)
    if isd:
        assert np.array_equal(m._a.ravel(order="F"), vals)
    else:
        assert np.array_equal(m._outer, outer) and np.array_equal(m._inner, inner) and np.array_equal(m._values, vals)
    return m


@pytest.mark.parametrize("name", ["A.txt", "B.txt"])
def test_reference_data_files(ref, name):
    if not os.path.exists(os.path.join(REF_DATA, name)):
        pytest.skip("reference data files not present")
    same_as_reference(ref, os.path.join(REF_DATA, name), np.complex128)


@pytest.mark.parametrize("cplx", [False, True])
@pytest.mark.parametrize("seed", range(4))
def test_generated_files_match_the_reference(ref, tmp_path, cplx, seed):
    rng = np.random.default_rng(seed)
    dt = np.complex128 if cplx else np.float64
    rows, cols = int(rng.integers(1, 40)), int(rng.integers(1, 40))
    a = rng.standard_normal((rows, cols)) * 10.0 ** rng.integers(-8, 8, (rows, cols))
    if cplx:
        a = a + 1j * rng.standard_normal((rows, cols))
    p = str(tmp_path / "d.txt")
    write_dense(p, a.astype(dt), cplx)
    same_as_reference(ref, p, dt)
    # sparse: unsorted entries, distinct positions, mixed whitespace
    nnz = int(rng.integers(1, rows * cols + 1))
    pos = rng.choice(rows * cols, size=nnz, replace=False)
    entries = [(int(q // cols), int(q % cols), complex(rng.standard_normal(), rng.standard_normal() if cplx else 0.0))
               for q in pos]
    p = str(tmp_path / "s.txt")
    write_sparse(p, rows, cols, entries, cplx, shuffle_ws=True)
    same_as_reference(ref, p, dt)


def test_large_sparse_file_parallel_chunks(ref, tmp_path, monkeypatch):
    """A file big enough for several parser threads (chunk boundaries fall inside entries)."""
    from pcsc_eigenvalue_solver_project_b200 import generators as gen
    g = 240
    rp, ci, v = gen.laplacian_2d(g)
    n = g * g
    rng = np.random.default_rng(1)
    order = rng.permutation(len(ci))
    rows_of = np.repeat(np.arange(n), np.diff(rp))
    p = str(tmp_path / "lap.txt")
    with open(p, "w") as f:
        f.write(f"sparse {n} {n} {len(ci)}\n")
        f.write("".join(f"{rows_of[k]} {ci[k]} {float(v[k])!r}\n" for k in order))
    assert os.path.getsize(p) > 2 << 20
    for threads in ("1", "3", "16"):
        monkeypatch.setenv("EIGSOL_B200_IO_THREADS", threads)
        m = same_as_reference(ref, p, np.float64)
    # and through the binary container
    b = str(tmp_path / "lap.bin")
    es.writeMatrixBinary(m, b)
    m2 = es.readMatrixBinary(b)
    assert np.array_equal(m2._outer, m._outer) and np.array_equal(m2._inner, m._inner) and np.array_equal(m2._values, m._values)


BAD = [
    ("", "Failed to read matrix storage type"),
    ("banded 3 3", "Unknown storage type: banded"),
    ("dense 3", "Failed to read matrix dimensions"),
    ("dense 0 3", "Matrix dimensions must be positive"),
    ("dense 2 2 1 2 3", "Failed to read scalar entry in dense matrix"),
    ("dense 2 2 1 2 x 4", "Failed to read scalar entry in dense matrix"),
    ("sparse 3 3", "Cannot read number of non-zero entries in the sparse matrix"),
    ("sparse 3 3 0", "number of non-zero entries must be positive in a sparse matrix"),
    ("sparse 3 3 2 0 0 1.0 1", "Error when trying to read indices in sparse matrix"),
    ("sparse 3 3 2 0 0 1.0 1 3 2.0", "Sparse indices out of range"),
    ("sparse 3 3 2 0 0 1.0 -1 1 2.0", "Sparse indices out of range"),
    ("sparse 3 3 2 0 0 1.0 1 1 abc", "Failed to read scalar entry in sparse matrix"),
    ("sparse 3 3 2 0 0 1.0 1 1", "Failed to read scalar entry in sparse matrix"),
]


@pytest.mark.parametrize("text,message", BAD)
def test_malformed_files_raise_the_reference_messages(tmp_path, text, message):
    p = str(tmp_path / "bad.txt")
    with open(p, "w") as f:
        f.write(text)
    with pytest.raises(RuntimeError) as e:
        es.readMatrixFromFile(p, np.float64)
    assert str(e.value) == message
    if have_reference():
        r = Reference()
        if hasattr(r.L, "ref_read_matrix_file"):
            with pytest.raises(RuntimeError) as e2:
                r.read_matrix_file(p, np.float64)
            assert str(e2.value) == message


def test_missing_file_and_complex_dense_message(tmp_path):
    with pytest.raises(RuntimeError, match="Impossible to open the file: /no/such/file"):
        es.readMatrixFromFile("/no/such/file")
    p = str(tmp_path / "c.txt")
    with open(p, "w") as f:
        f.write("dense 1 2 1 2 3")
    with pytest.raises(RuntimeError, match="Failed to read complex entry in dense matrix"):
        es.readMatrixFromFile(p, np.complex128)


def test_binary_container_round_trip_and_rejects_garbage(tmp_path):
    rng = np.random.default_rng(3)
    a = rng.standard_normal((7, 5)) + 1j * rng.standard_normal((7, 5))
    p = str(tmp_path / "d.bin")
    es.writeMatrixBinary(es.Matrix(a), p)
    m = es.readMatrixBinary(p)
    assert m.isDense() and np.array_equal(m._a, a)
    with open(p, "r+b") as f:
        f.write(b"NOTMAGIC")
    with pytest.raises(RuntimeError, match="Not an EIGSOLB2 matrix file"):
        es.readMatrixBinary(p)


def _patch(path, offset, fmt, value):
    import struct
    with open(path, "r+b") as f:
        f.seek(offset)
        f.write(struct.pack(fmt, value))


def test_binary_container_rejects_corrupt_headers_and_truncation(tmp_path):
    """The EIGSOLB2 header is untrusted: wrapped offsets, a short file, a dense flag outside {0,1} and column pointers
    that disagree with the header must all come back as the IO error, never as an out-of-bounds read.
    Header layout: magic[8], version i32, dtype i32, dense i32, pad i32, rows i64, cols i64, nnz i64."""
    import shutil
    import scipy.sparse as sp
    g = 6
    rp, ci, v = gen.laplacian_2d(g)
    a = sp.csr_matrix((v, ci, rp), shape=(g * g, g * g)).tocsc()
    good = str(tmp_path / "good.bin")
    es.writeMatrixBinary(es.Matrix(a), good)
    m = es.readMatrixBinary(good)
    assert not m.isDense() and np.array_equal(m._values, a.data)
    size = os.path.getsize(good)
    cases = {
        "nnz_huge": (40, "<q", 2**62),           # s * nnz wraps in 64-bit arithmetic
        "nnz_above_int32": (40, "<q", 2**31 + 5),
        "cols_huge": (32, "<q", 2**40),
        "rows_above_int32": (24, "<q", 2**31),
        "dense_flag": (16, "<i", 7),
        "nnz_disagrees_with_outer": (40, "<q", int(a.nnz) - 1),
    }
    for name, (off, fmt, val) in cases.items():
        p = str(tmp_path / f"{name}.bin")
        shutil.copy(good, p)
        _patch(p, off, fmt, val)
        with pytest.raises(RuntimeError, match="EIGSOLB2"):
            es.readMatrixBinary(p)
    # outer[0] != 0
    p = str(tmp_path / "outer0.bin")
    shutil.copy(good, p)
    _patch(p, 64, "<i", 1)
    with pytest.raises(RuntimeError, match="EIGSOLB2"):
        es.readMatrixBinary(p)
    # truncation: 100 bytes short loses payload of the values array; 1 byte short of the payload end too
    payload_end = 64 + ((4 * (g * g + 1) + 63) // 64) * 64 + ((4 * a.nnz + 63) // 64) * 64 + 8 * a.nnz
    for cut in (size - 100, payload_end - 1):
        p = str(tmp_path / f"cut{cut}.bin")
        with open(good, "rb") as f, open(p, "wb") as o:
            o.write(f.read(cut))
        with pytest.raises(RuntimeError, match="EIGSOLB2"):
            es.readMatrixBinary(p)
    # the alignment padding behind the LAST array is optional
    p = str(tmp_path / "nopad.bin")
    with open(good, "rb") as f, open(p, "wb") as o:
        o.write(f.read(payload_end))
    m = es.readMatrixBinary(p)
    assert np.array_equal(m._values, a.data)
    # dense file whose nnz is not rows * cols
    d = str(tmp_path / "dense.bin")
    es.writeMatrixBinary(es.Matrix(np.arange(12.0).reshape(3, 4)), d)
    _patch(d, 40, "<q", 11)
    with pytest.raises(RuntimeError, match="EIGSOLB2"):
        es.readMatrixBinary(d)


# ---- the C++ drop-in header (eigsol/src/reader/file_matrix_reader.hpp) ---------------------------------------
DUMP = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "build", "dropin", "bin", "reader_dump")


def cxx_read(path, kind, tmp_path):
    """EigSol::readMatrixFromFile<Scalar>(path) through the drop-in header; returns (rc, stderr, dump or None)."""
    import subprocess
    if not os.path.exists(DUMP):
        pytest.skip("build/dropin/bin/reader_dump not built (needs the reference checkout for Eigen)")
    out = str(tmp_path / "dump.bin")
    p = subprocess.run([DUMP, kind, path, out], capture_output=True, text=True, timeout=120)
    if p.returncode != 0:
        return p.returncode, p.stderr.strip(), None
    raw = open(out, "rb").read()
    isd, rows, cols, nnz = (int(x) for x in np.frombuffer(raw[:32], dtype=np.int64))
    off = 32
    outer = inner = None
    if not isd:
        outer = np.frombuffer(raw, dtype=np.int32, count=cols + 1, offset=off)
        off += 4 * (cols + 1)
        inner = np.frombuffer(raw, dtype=np.int32, count=nnz, offset=off)
        off += 4 * nnz
    vals = np.frombuffer(raw, dtype=np.float64, offset=off)
    if kind == "z":
        vals = vals.view(np.complex128)
    assert len(vals) == nnz
    return 0, "", (bool(isd), rows, cols, outer, inner, vals)


@pytest.mark.parametrize("cplx", [False, True])
def test_cxx_header_returns_the_same_matrix(tmp_path, cplx):
    rng = np.random.default_rng(11 + cplx)
    dt = np.complex128 if cplx else np.float64
    kind = "z" if cplx else "d"
    a = rng.standard_normal((9, 6)) + (1j * rng.standard_normal((9, 6)) if cplx else 0)
    p = str(tmp_path / "d.txt")
    write_dense(p, a.astype(dt), cplx)
    rc, err, got = cxx_read(p, kind, tmp_path)
    assert rc == 0, err
    m = es.readMatrixFromFile(p, dt)
    assert got[0] and got[1:3] == (9, 6) and np.array_equal(got[5], m._a.ravel(order="F"))
    pos = rng.choice(54, size=30, replace=False)
    entries = [(int(q // 6), int(q % 6), complex(rng.standard_normal(), rng.standard_normal() if cplx else 0.0)) for q in pos]
    p = str(tmp_path / "s.txt")
    write_sparse(p, 9, 6, entries, cplx, shuffle_ws=True)
    rc, err, got = cxx_read(p, kind, tmp_path)
    assert rc == 0, err
    m = es.readMatrixFromFile(p, dt)
    assert not got[0] and got[1:3] == (9, 6)
    assert np.array_equal(got[3], m._outer) and np.array_equal(got[4], m._inner) and np.array_equal(got[5], m._values)


def test_cxx_header_generic_scalar_path(tmp_path):
    """float goes through the stream loop of the header (the C ABI parses double / complex<double> only)."""
    entries = [(2, 1, 0.1), (0, 0, 1.5), (1, 2, -3.25e-3), (0, 2, 7.0)]
    p = str(tmp_path / "s.txt")
    write_sparse(p, 3, 3, entries, False)
    rc, err, got = cxx_read(p, "f", tmp_path)
    assert rc == 0, err
    m = es.readMatrixFromFile(p, np.float64)
    assert np.array_equal(got[3], m._outer) and np.array_equal(got[4], m._inner)
    assert np.array_equal(got[5], m._values.astype(np.float32).astype(np.float64))
    p = str(tmp_path / "d.txt")
    write_dense(p, np.array([[0.1, 2.0], [3.0, -0.7]]), False)
    rc, err, got = cxx_read(p, "f", tmp_path)
    assert rc == 0 and got[0]
    assert np.array_equal(got[5], np.array([0.1, 3.0, 2.0, -0.7], dtype=np.float32).astype(np.float64))


@pytest.mark.parametrize("kind", ["d", "f"])
@pytest.mark.parametrize("text,message", BAD)
def test_cxx_header_raises_the_reference_messages(tmp_path, kind, text, message):
    p = str(tmp_path / "bad.txt")
    with open(p, "w") as f:
        f.write(text)
    rc, err, _ = cxx_read(p, kind, tmp_path)
    assert rc == 3 and err == message


def test_duplicate_entries_are_rejected_by_every_reader_path(tmp_path):
    """A repeated (row, column) pair: the reference's SparseMatrix::insert forbids it.  The fast reader (double /
    complex<double>, also behind the Python wrapper) and the drop-in header's stream path (other scalars, here float)
    answer with the same error."""
    p = str(tmp_path / "dup.txt")
    open(p, "w").write("sparse\n3 3 4\n0 0 1.0\n1 1 2.0\n0 0 3.0\n2 2 4.0\n")
    with pytest.raises(RuntimeError, match="Duplicate \\(row, column\\)"):
        es.readMatrixFromFile(p, np.float64)
    for kind in ("d", "f"):
        rc, err, _ = cxx_read(p, kind, tmp_path)
        assert rc == 3 and "Duplicate (row, column)" in err, (kind, rc, err)
