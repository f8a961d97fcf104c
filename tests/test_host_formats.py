"""Stand-alone C++ host (host/): LIBSVM / ARFF / model parsers, model writer and the plssvm-train / plssvm-predict command lines.

CPU part: the host's parsers against the REFERENCE's own parsers (src/plssvm/parameter.cpp, compiled unmodified into oracle/_ref)
on files written here — same arrays bit for bit, same quirks (0-based indices, width = max index + 1, sign() on labels, comment and
blank lines), same error texts as the reference's tests assert (tests/parameter_test.cpp:135-630); `--help`; loud failure without a GPU.
GPU part: train -> model file -> predict round trip through the executables, model text identical to the reference's write_model
for the same alpha (src/plssvm/csvm.cpp:60-204).
"""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from oracle import port, ref
from plssvm_sparse_fp22_b200 import datagen

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "host", "bin")
HOSTLIB = os.path.join(BIN, "libplssvm_b200_host.so")


@pytest.fixture(scope="module")
def host():
    if not os.path.exists(HOSTLIB):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "host")], stdout=subprocess.DEVNULL)
    L = C.CDLL(HOSTLIB)
    vp, i, d, l = C.c_void_p, C.c_int, C.c_double, C.c_long
    pl, pi, pd = C.POINTER(l), C.POINTER(i), C.POINTER(d)
    L.b200host_last_error.restype = C.c_char_p
    L.b200host_parse_data_file.argtypes = [i, C.c_char_p, pl, pl, pl, pi, vp, vp, pd]
    L.b200host_parse_model_file.argtypes = [i, C.c_char_p, pi, pi, pd, pd, pd, pl, pl, vp, vp, vp]
    L.b200host_format_shortest.argtypes = [i, d, C.c_char_p, i]
    L.b200host_fp22_encode.argtypes = [C.c_float]
    L.b200host_fp22_encode.restype = C.c_uint
    return L


def host_parse(L, path, dt=np.float64):
    n, F, nnz, has, g = C.c_long(0), C.c_long(0), C.c_long(0), C.c_int(0), C.c_double(0)
    code = 0 if dt == np.float32 else 1
    rc = L.b200host_parse_data_file(code, path.encode(), C.byref(n), C.byref(F), C.byref(nnz), C.byref(has), None, None, C.byref(g))
    if rc != 0:
        return rc, L.b200host_last_error().decode()
    X = np.empty((n.value, F.value), dt)
    y = np.empty(n.value, dt)
    L.b200host_parse_data_file(code, path.encode(), C.byref(n), C.byref(F), C.byref(nnz), C.byref(has), X.ctypes.data, y.ctypes.data, C.byref(g))
    return 0, (X, y if has.value else None, nnz.value, g.value)


def ref_parse(path, dt=np.float64):
    L = ref.lib()
    L.oref_parse_data_file.argtypes = [C.c_int, C.c_char_p, C.POINTER(C.c_long), C.POINTER(C.c_long), C.POINTER(C.c_int), C.c_void_p, C.c_void_p]
    n, F, has = C.c_long(0), C.c_long(0), C.c_int(0)
    code = 0 if dt == np.float32 else 1
    assert L.oref_parse_data_file(code, path.encode(), C.byref(n), C.byref(F), C.byref(has), None, None) == 0
    X = np.empty((n.value, F.value), dt)
    y = np.empty(n.value, dt)
    L.oref_parse_data_file(code, path.encode(), C.byref(n), C.byref(F), C.byref(has), X.ctypes.data, y.ctypes.data)
    return X, y if has.value else None


LIBSVM_TEXT = """# comment line
1 0:1.5 3:-2.25e-1 7:0.000012

-1 1:3 2:4.5
  -1 0:0.1 7:1e10 3:0
# another comment
1 2:-7
"""

ARFF_TEXT = """% comment
@RELATION test
@ATTRIBUTE a NUMERIC
@ATTRIBUTE b numeric
@ATTRIBUTE c NUMERIC
@ATTRIBUTE class NUMERIC
@DATA
1.5,0,-2,1
{0 3.25,3 -1}
% in between
0,0,1e-3,-1
{1 7.0,2 0.5,3 1}
"""


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref not built")
@pytest.mark.parametrize("dt", [np.float64, np.float32])
def test_parsers_match_reference(host, tmp_path, dt):
    cases = {"a.libsvm": LIBSVM_TEXT, "b.arff": ARFF_TEXT, "nolabel.libsvm": "0:1 2:3\n1:4.5\n"}
    X, y = datagen.make_dense(40, 9, seed=2, problem="planes", dtype=np.float64)
    X[np.abs(X) < 0.3] = 0.0
    datagen.write_libsvm(str(tmp_path / "gen.libsvm"), X, y)
    cases["gen.libsvm"] = None
    for name, text in cases.items():
        p = str(tmp_path / name)
        if text is not None:
            open(p, "w").write(text)
        rc, res = host_parse(host, p, dt)
        assert rc == 0, res
        Xh, yh, nnz, gamma = res
        Xr, yr = ref_parse(p, dt)
        assert Xh.shape == Xr.shape and np.array_equal(Xh, Xr), name
        assert (yh is None) == (yr is None)
        if yh is not None:
            assert np.array_equal(yh, yr)
        assert nnz == int(np.count_nonzero(Xr))
        assert gamma == pytest.approx(float(dt(1.0) / dt(Xr.shape[1])))


@pytest.mark.parametrize("text,msg", [
    ("1 0:1 1:abc\n", "Can't convert 'abc' to a value of type double!"),
    ("1 x:1\n", "Can't convert 'x' to a value of type unsigned long!"),
    ("# only a comment\n\n", "Can't parse file: no data points are given!"),
])
def test_libsvm_errors(host, tmp_path, text, msg):
    p = str(tmp_path / "bad.libsvm")
    open(p, "w").write(text)
    rc, err = host_parse(host, p)
    assert rc == 2 and msg in err, err


@pytest.mark.parametrize("text,msg", [
    ("@RELATION r\n@ATTRIBUTE a STRING\n@DATA\n1\n", "Can only use NUMERIC features, but '@ATTRIBUTE a STRING' was given!"),
    ("@RELATION r\n@ATTRIBUTE class NUMERIC\n@ATTRIBUTE b NUMERIC\n@DATA\n1,2\n", "Only the last ATTRIBUTE may be CLASS!"),
    ("@RELATION r\n@DATA\n1,2\n", "Can't parse file: no ATTRIBUTES are defined!"),
    ("@RELATION r\n@ATTRIBUTE a NUMERIC\n@DATA\n", "Can't parse file: no data points are given or @DATA is missing!"),
    ("@RELATION r\n@ATTRIBUTE a NUMERIC\n@ATTRIBUTE b NUMERIC\n@DATA\n1,2\n@ATTRIBUTE c NUMERIC\n", "Read @ inside data section!: '@ATTRIBUTE c NUMERIC'"),
    ("@RELATION r\n@ATTRIBUTE a NUMERIC\n@ATTRIBUTE b NUMERIC\n@DATA\n{0 1,1 2\n", "Missing closing '}' for sparse data point 0 description!"),
    ("@RELATION r\n@ATTRIBUTE a NUMERIC\n@ATTRIBUTE b NUMERIC\n@DATA\n{5 1}\n", "Too many features given! Trying to add feature at position 5 but max position is 1!"),
    ("@RELATION r\n@ATTRIBUTE a NUMERIC\n@ATTRIBUTE class NUMERIC\n@DATA\n{0 1}\n", "Missing label for data point 0!"),
    ("@RELATION r\n@ATTRIBUTE a NUMERIC\n@ATTRIBUTE b NUMERIC\n@DATA\n1\n", "Invalid number of features/labels! Found 0 but should be 1!"),
    ("@RELATION r\n@ATTRIBUTE a NUMERIC\n@ATTRIBUTE b NUMERIC\n@DATA\n1,2,3\n", "Too many features! Superfluous ',3' for data point 0!"),
])
def test_arff_errors(host, tmp_path, text, msg):
    p = str(tmp_path / "bad.arff")
    open(p, "w").write(text)
    rc, err = host_parse(host, p)
    assert rc == 2 and msg in err, err


def test_missing_file(host, tmp_path):
    rc, err = host_parse(host, str(tmp_path / "nope.libsvm"))
    assert rc == 3 and "Couldn't find file" in err


MODEL_HEADER = "svm_type c_svc\nkernel_type rbf\ngamma 0.25\nnr_class 2\ntotal_sv 3\nrho 0.5\nlabel 1 -1\nnr_sv 2 1\nSV\n"
MODEL_SVS = "0.125 0:1.000000e+00 2:-5.000000e-01 \n-1e-05 1:2.500000e+00 \n3 0:1.000000e-03 3:4.000000e+00 \n"


def host_parse_model(L, path):
    k, dg, n, F = C.c_int(0), C.c_int(0), C.c_long(0), C.c_long(0)
    g, c0, rho = C.c_double(0), C.c_double(0), C.c_double(0)
    rc = L.b200host_parse_model_file(1, path.encode(), C.byref(k), C.byref(dg), C.byref(g), C.byref(c0), C.byref(rho), C.byref(n), C.byref(F), None, None, None)
    if rc != 0:
        return rc, L.b200host_last_error().decode()
    X = np.empty((n.value, F.value))
    a = np.empty(n.value)
    y = np.empty(n.value)
    L.b200host_parse_model_file(1, path.encode(), C.byref(k), C.byref(dg), C.byref(g), C.byref(c0), C.byref(rho), C.byref(n), C.byref(F), X.ctypes.data, a.ctypes.data,
                                y.ctypes.data)
    return 0, dict(kernel=k.value, degree=dg.value, gamma=g.value, coef0=c0.value, rho=rho.value, X=X, alpha=a, y=y)


def test_model_parser(host, tmp_path):
    p = str(tmp_path / "m.model")
    open(p, "w").write(MODEL_HEADER + MODEL_SVS)
    rc, m = host_parse_model(host, p)
    assert rc == 0, m
    assert m["kernel"] == 2 and m["gamma"] == 0.25 and m["rho"] == 0.5
    assert np.array_equal(m["alpha"], [0.125, -1e-05, 3.0])
    assert np.array_equal(m["y"], [1, 1, -1])
    assert np.array_equal(m["X"], [[1, 0, -0.5, 0], [0, 2.5, 0, 0], [1e-3, 0, 0, 4]])
    for bad, msg in [
        (MODEL_HEADER.replace("c_svc", "nu_svc"), "Can only use c_svc as svm_type, but 'nu_svc' was given!"),
        (MODEL_HEADER.replace("rbf", "sigmoid"), "Unrecognized kernel type 'sigmoid'!"),
        (MODEL_HEADER.replace("nr_class 2", "nr_class 3"), "Can only use 2 classes, but 3 were given!"),
        (MODEL_HEADER.replace("rho 0.5\n", ""), "Missing rho value!"),
        (MODEL_HEADER.replace("nr_sv 2 1", "nr_sv 2 2"), "doesn't add up to the total number: 2 + 2 != 3!"),
        (MODEL_HEADER.replace("label 1 -1", "label 1 2"), "Only the labels 1 and -1 are allowed"),
        (MODEL_HEADER.replace("SV\n", "foo 1\nSV\n"), "Unrecognized header entry 'foo 1'! Maybe SV is missing?"),
    ]:
        open(p, "w").write(bad + MODEL_SVS)
        rc, err = host_parse_model(host, p)
        assert rc == 2 and msg in err, err


def test_number_formatting_like_fmt(host):
    buf = C.create_string_buffer(64)
    for v, want in [(1.0, "1"), (0.5, "0.5"), (-0.0625, "-0.0625"), (1e-5, "1e-05"), (0.0001, "0.0001"), (123456789.0, "123456789"), (1e16, "1e+16"),
                    (0.1, "0.1"), (2.5e-7, "2.5e-07"), (1234.5678, "1234.5678"), (0.0, "0"), (1e15, "1000000000000000")]:
        assert host.b200host_format_shortest(1, v, buf, 64) == 0
        assert buf.value.decode() == want, (v, buf.value)
    assert host.b200host_format_shortest(0, 0.1, buf, 64) == 0 and buf.value.decode() == "0.1"  # float shortest, not 0.100000001
    for x in (0.1, -3.7, 1e-20, 65504.0, 0.33333334):
        assert host.b200host_fp22_encode(x) == port.fp22_encode(x)


def test_cli_help_and_loud_failure_without_gpu(host, tmp_path):
    for exe in ("plssvm-train", "plssvm-predict"):
        r = subprocess.run([os.path.join(BIN, exe), "--help"], capture_output=True, text=True)
        assert r.returncode == 0 and "Usage:" in r.stdout and "--backend" in r.stdout
    r = subprocess.run([os.path.join(BIN, "plssvm-train")], capture_output=True, text=True)
    assert r.returncode == 1 and "Error missing input file!" in r.stderr
    import torch
    if not torch.cuda.is_available():
        p = str(tmp_path / "t.libsvm")
        open(p, "w").write(LIBSVM_TEXT)
        r = subprocess.run([os.path.join(BIN, "plssvm-train"), "-q", p, str(tmp_path / "t.model")], capture_output=True, text=True)
        assert r.returncode == 1 and "no CPU fallback" in r.stderr and "backend_exception" in r.stderr
        r = subprocess.run([os.path.join(BIN, "plssvm-train"), "-q", "-b", "openmp", p], capture_output=True, text=True)
        assert r.returncode == 1 and "No openmp backend available!" in r.stderr
        # cxxopts takes the positionals by name as well (parameter_train.cpp:60-61): the file is parsed before the device is asked for
        r = subprocess.run([os.path.join(BIN, "plssvm-train"), "--model", str(tmp_path / "n.model"), "--input", p], capture_output=True, text=True)
        assert r.returncode == 1 and "no CPU fallback" in r.stderr and f"input file (data set): '{p}'" in r.stdout and "n.model'" in r.stdout
    r = subprocess.run([os.path.join(BIN, "plssvm-predict"), "--model", "m"], capture_output=True, text=True)
    assert r.returncode == 1 and "Error missing test file!" in r.stderr
    r = subprocess.run([os.path.join(BIN, "plssvm-predict"), "--test=t"], capture_output=True, text=True)
    assert r.returncode == 1 and "Error missing model file!" in r.stderr
    r = subprocess.run([os.path.join(BIN, "plssvm-train"), "--nope"], capture_output=True, text=True)
    assert r.returncode == 1 and "Option 'nope' does not exist" in r.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("kernel,flags", [("linear", ["-t", "0"]), ("polynomial", ["-t", "1", "-d", "2", "-r", "0.5"]), ("rbf", ["--kernel_type=rbf", "-g", "0.02"])])
def test_cli_train_predict_round_trip(host, tmp_path, kernel, flags):
    X, y = datagen.make_dense(300, 40, seed=4, problem="planes", dtype=np.float64)
    Xt, yt = datagen.make_dense(120, 40, seed=5, problem="planes", dtype=np.float64)
    train, test, model, pred = (str(tmp_path / n) for n in ("train.libsvm", "test.libsvm", "train.model", "out.predict"))
    datagen.write_libsvm(train, X, y)
    datagen.write_libsvm(test, Xt, yt)
    r = subprocess.run([os.path.join(BIN, "plssvm-train"), *flags, "-e", "1e-6", train, model], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "task: training" in r.stdout and "Finished after" in r.stdout
    rc, m = host_parse_model(host, model)
    assert rc == 0, m
    kw = dict(degree=2, coef0=0.5) if kernel == "polynomial" else {}
    gamma = 0.02 if kernel == "rbf" else 1.0 / 40
    a_o, rho_o, _ = port.learn(X, y, kernel, gamma=gamma, eps=1e-6, use_rows=True, **kw)
    order = np.concatenate([np.nonzero(y > 0)[0], np.nonzero(y < 0)[0]])  # +1 rows first, then -1 rows
    assert np.linalg.norm(m["alpha"] - a_o[order]) / np.linalg.norm(a_o) < 1e-6
    assert abs(m["rho"] - rho_o) < 1e-5 * max(1.0, abs(rho_o))  # rho is a cancellation of large terms
    assert np.allclose(m["X"], X[order], rtol=1e-6, atol=0)
    # the model text equals what the reference's own write_model produces for the same alpha / data / rho
    if ref.available():
        L = ref.lib()
        L.oref_write_model.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_void_p, C.c_void_p, C.c_void_p, C.c_long, C.c_long, C.c_double, C.c_char_p]
        alpha_file_order = np.empty_like(a_o)
        alpha_file_order[order] = m["alpha"]
        ref.set_num_threads(1)
        refmodel = str(tmp_path / "ref.model")
        assert L.oref_write_model(1, port.KERNELS[kernel], kw.get("degree", 3), gamma, kw.get("coef0", 0.0), X.ctypes.data, y.ctypes.data,
                                  alpha_file_order.ctypes.data, 300, 40, m["rho"], refmodel.encode()) == 0
        assert open(refmodel).read() == open(model).read()
    r = subprocess.run([os.path.join(BIN, "plssvm-predict"), test, model, pred], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    labels = np.array([float(v) for v in open(pred).read().split("\n")])
    assert not open(pred).read().endswith("\n")
    dec_o = port.predict(X, a_o, rho_o, Xt, kernel, gamma=gamma, **kw)
    safe = np.abs(dec_o) > 1e-6
    assert np.array_equal(labels[safe], port.sign(dec_o)[safe])
    acc = float((labels * yt > 0).mean())
    assert f"({int((labels * yt > 0).sum())}/120) (classification)" in r.stdout and acc > 0.5


@pytest.mark.gpu
def test_cli_sparse_fp22_round_trip(host, tmp_path):
    rp, ci, v, y = datagen.make_sparse_csr(500, 3000, 0.01, seed=8, dtype=np.float32)
    X = datagen.csr_to_dense(rp, ci, v, 3000)
    X[-1, -1] = 0.5  # make sure the last feature index appears so that the parsed width is 3000
    train, model, pred = (str(tmp_path / n) for n in ("s.libsvm", "s.model", "s.predict"))
    datagen.write_libsvm(train, X, y)
    # gamma = 0.05: with the default 1/F every kernel value is ~1 and the float solve is ill-conditioned (rho is a difference of
    # sums ~500 of alpha; all float paths disagree with the double oracle there, so labels near 0 are not comparable)
    r = subprocess.run([os.path.join(BIN, "plssvm-train"), "-t", "2", "-g", "0.05", "-e", "1e-4", "--fp22", train, model], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = subprocess.run([os.path.join(BIN, "plssvm-predict"), "--fp22", "-q", train, model, pred], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    labels = np.array([float(s) for s in open(pred).read().split("\n")])
    Xq = datagen.fp22_round(X).astype(np.float64)
    a_o, rho_o, _ = port.learn(Xq, y.astype(np.float64), "rbf", gamma=0.05, eps=1e-4, use_rows=True)
    dec_o = port.predict(Xq, a_o, rho_o, Xq, "rbf", gamma=0.05)
    safe = np.abs(dec_o) > 1e-3 * np.abs(dec_o).max()
    assert np.array_equal(labels[safe], port.sign(dec_o)[safe])


def test_parallel_libsvm_parser_is_order_preserving(host, tmp_path):
    """The LIBSVM parser splits the lines over an OpenMP team (like the reference, src/plssvm/parameter.cpp:55-99): the result
    must not depend on the team size, and of several malformed lines the first one in file order is the one reported."""
    rng = np.random.default_rng(12)
    n, F = 3000, 60
    X = np.where(rng.random((n, F)) < 0.2, rng.standard_normal((n, F)), 0.0)
    X[:, -1] = rng.standard_normal(n)  # the width of the file is F
    y = np.where(rng.random(n) < 0.5, 1.0, -1.0)
    path = str(tmp_path / "big.libsvm")
    datagen.write_libsvm(path, X, y)
    results = []
    for threads in (1, 2, 7, 64):
        host.b200host_set_threads(threads)
        rc, out = host_parse(host, path)
        assert rc == 0, out
        results.append(out)
    X0, y0, nnz0, g0 = results[0]
    assert np.array_equal(X0, X) and np.array_equal(y0, y) and nnz0 == int((X != 0).sum())
    for Xk, yk, nnzk, gk in results[1:]:
        assert np.array_equal(Xk, X0) and np.array_equal(yk, y0) and nnzk == nnz0 and gk == g0
    # two malformed lines: line 500 (bad value) comes before line 2500 (bad index)
    lines = open(path).read().split("\n")
    lines[500] = "1 0:0.5 3:abc"
    lines[2500] = "1 x:0.5"
    bad = str(tmp_path / "bad.libsvm")
    open(bad, "w").write("\n".join(lines))
    msgs = set()
    for threads in (1, 5, 64):
        host.b200host_set_threads(threads)
        rc, msg = host_parse(host, bad)
        assert rc != 0
        msgs.add(msg)
    assert msgs == {"Can't convert 'abc' to a value of type double!"}
    host.b200host_set_threads(os.cpu_count() or 1)


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref not built")
@pytest.mark.parametrize("dt", [np.float64, np.float32])
@pytest.mark.parametrize("kernel", ["linear", "polynomial", "rbf"])
def test_model_writer_matches_reference_text(host, tmp_path, kernel, dt):
    """The model text of the host writer (rows formatted by an OpenMP team, std::to_chars) against the reference's own write_model
    (src/plssvm/csvm.cpp:60-204, run with one thread so that its order inside a class is the file order): byte-identical for dense
    and CSR storage, any team size, sizes that span several batches of the writer, values that print with 2- and 3-digit exponents."""
    rng = np.random.default_rng(11)
    for n, F in ((7, 5), (300, 40), (70001, 3)):
        X = rng.uniform(-1, 1, (n, F))
        X[rng.random((n, F)) < 0.4] = 0.0
        X[0, 0], X[1, F - 1] = 1.5e-30, -7.25e+33  # exponent widths
        if dt == np.float64:
            X[2, 1] = 3.0e-200
        X = X.astype(dt)
        y = np.where(rng.random(n) < 0.5, 1.0, -1.0).astype(dt)
        y[:2] = (1.0, -1.0)
        alpha = rng.standard_normal(n).astype(dt)
        rho, gamma, coef0, degree = -0.3712, 0.02, 0.5, 2
        code = 1 if dt == np.float64 else 0
        L = ref.lib()
        L.oref_write_model.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_void_p, C.c_void_p, C.c_void_p, C.c_long, C.c_long, C.c_double, C.c_char_p]
        host.b200host_write_model.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_void_p, C.c_void_p, C.c_void_p, C.c_long, C.c_long, C.c_double,
                                              C.c_char_p, C.c_int]
        ref.set_num_threads(1)
        want_path = str(tmp_path / "ref.model")
        assert L.oref_write_model(code, port.KERNELS[kernel], degree, gamma, coef0, X.ctypes.data, y.ctypes.data, alpha.ctypes.data, n, F, rho, want_path.encode()) == 0
        want = open(want_path, "rb").read()
        try:
            for threads in (1, 3, os.cpu_count() or 1):
                host.b200host_set_threads(threads)
                for as_csr in (0, 1):
                    got_path = str(tmp_path / f"host_{threads}_{as_csr}.model")
                    rc = host.b200host_write_model(code, port.KERNELS[kernel], degree, gamma, coef0, X.ctypes.data, y.ctypes.data, alpha.ctypes.data, n, F, rho,
                                                   got_path.encode(), as_csr)
                    assert rc == 0, host.b200host_last_error()
                    assert open(got_path, "rb").read() == want, (n, F, threads, as_csr)
        finally:
            host.b200host_set_threads(os.cpu_count() or 1)


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref not built")
def test_parallel_arff_parser_matches_reference(host, tmp_path):
    """The ARFF data section is parsed by an OpenMP team straight into CSR (no dense row per line): dense and sparse lines mixed,
    repeated indices inside a sparse line (the later entry wins, as in the reference's dense row), explicit zeros, comments in
    between — equal to the reference's parser for every team size; of two malformed lines the first in file order is reported."""
    rng = np.random.default_rng(21)
    n, F = 2500, 37
    lines = ["% generated", "@RELATION big"] + [f"@ATTRIBUTE f{j} NUMERIC" for j in range(F)] + ["@ATTRIBUTE class NUMERIC", "@DATA"]
    for i in range(n):
        label = 1 if rng.random() < 0.5 else -1
        if i % 3 == 0:
            vals = np.where(rng.random(F) < 0.5, np.round(rng.standard_normal(F), 6), 0.0)
            lines.append(",".join(repr(float(v)) for v in vals) + f",{label}")
        else:
            idx = rng.integers(0, F, rng.integers(1, 9))  # unsorted, may repeat
            ent = [f"{j} {0.0 if rng.random() < 0.1 else round(float(rng.standard_normal()), 6)!r}" for j in idx]
            lines.append("{" + ",".join(ent + [f"{F} {label}"]) + "}")  # the class entry comes last (parameter.cpp:270-276 reads it to the end of the line)
        if i % 400 == 7:
            lines.append("% a comment inside the data section")
    path = str(tmp_path / "big.arff")
    open(path, "w").write("\n".join(lines) + "\n")
    Xr, yr = ref_parse(path, np.float64)
    assert Xr.shape == (n, F)
    for threads in (1, 2, 7, 64):
        host.b200host_set_threads(threads)
        rc, out = host_parse(host, path)
        assert rc == 0, out
        Xh, yh, nnz, gamma = out
        assert np.array_equal(Xh, Xr) and np.array_equal(yh, yr) and nnz == int(np.count_nonzero(Xr))
    data0 = lines.index("@DATA") + 1
    bad_lines = list(lines)
    bad_lines[data0 + 300] = "{0 1.5,3 abc," + f"{F} 1" + "}"
    bad_lines[data0 + 2000] = "@ATTRIBUTE late NUMERIC"
    bad = str(tmp_path / "bad.arff")
    open(bad, "w").write("\n".join(bad_lines) + "\n")
    msgs = set()
    for threads in (1, 5, 64):
        host.b200host_set_threads(threads)
        rc, msg = host_parse(host, bad)
        assert rc != 0
        msgs.add(msg)
    assert msgs == {"Can't convert 'abc' to a value of type double!"}
    host.b200host_set_threads(os.cpu_count() or 1)
