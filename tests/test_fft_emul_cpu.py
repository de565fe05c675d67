"""CPU: the shared-memory line FFTs of pyscf_isdf_b200/csrc/fft_core.cuh, compiled with g++ and run
with a CTA emulated by threads + a barrier (tests/emul/), against numpy.fft -- the pin the reference
uses for its own FFT (pyscf/pbc/tools/test/test_pbc.py:185-217).  Covers the routines the GPU kernels
call (Stockham radix 2/3/4/5/7, direct DFT for other prime factors, Rader for prime lengths) and the
opt-in fused Rader path (ISDF_FFT_FUSED=1: permutation / kernel multiply / un-permutation folded into
the sub-FFT stages, radix 10), in both memory layouts the kernels use.  Tolerance 1e-12 relative."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, 'emul', 'fft_emul.cpp')
OUT = os.path.join(HERE, 'emul', '_build', 'libfft_emul.so')
DEPS = [SRC, os.path.join(HERE, 'emul', 'cuda_shim.h'),
        os.path.join(HERE, '..', 'pyscf_isdf_b200', 'csrc', 'fft_core.cuh')]


@pytest.fixture(scope='module')
def emul():
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    if not os.path.exists(OUT) or any(os.path.getmtime(d) > os.path.getmtime(OUT) for d in DEPS):
        subprocess.check_call(['g++', '-std=c++20', '-O1', '-pthread', '-fPIC', '-shared', '-w', '-o', OUT, SRC])
    lib = ctypes.CDLL(OUT)
    lib.emul_fft.restype = ctypes.c_int
    lib.emul_fft.argtypes = [ctypes.c_int] * 6 + [ctypes.c_void_p, ctypes.c_void_p, ctypes.POINTER(ctypes.c_int)]
    return lib


def _run(lib, x, sign, fused, line_fast, nthreads):
    nlines, n = x.shape
    xin = np.ascontiguousarray(x, dtype=np.complex128)
    out = np.empty_like(xin)
    used = ctypes.c_int(0)
    rc = lib.emul_fft(n, nlines, sign, fused, line_fast, nthreads, xin.ctypes.data, out.ctypes.data, ctypes.byref(used))
    assert rc == 0
    return out, used.value


def _ref(x, sign):
    return np.fft.fft(x, axis=1) if sign < 0 else np.fft.ifft(x, axis=1) * x.shape[1]


# mesh sizes of the BASELINE configs (35, 71, 96, 108), small primes with Rader sub-lengths of every
# shape (11 -> 10 one fused stage, 13 -> 12, 23 -> 2*11 direct-DFT factor, 31 -> 30, 41 -> 40, 101 -> 100),
# a big prime factor (2*13), and plain composites
LENGTHS = [1, 2, 3, 4, 5, 7, 8, 9, 10, 11, 12, 13, 23, 26, 31, 35, 36, 41, 48, 64, 71, 96, 97, 100, 101, 108, 128]


@pytest.mark.parametrize('line_fast', [0, 1])
@pytest.mark.parametrize('fused', [0, 1])
def test_line_ffts_match_numpy(emul, fused, line_fast):
    rng = np.random.default_rng(11 + 2 * fused + line_fast)
    took_fused = {}
    for n in LENGTHS:
        nlines = 1 if n == 128 else 5
        x = rng.standard_normal((nlines, n)) + 1j * rng.standard_normal((nlines, n))
        for sign in (-1, 1):
            got, used = _run(emul, x, sign, fused, line_fast, nthreads=7)
            ref = _ref(x, sign)
            err = abs(got - ref).max() / abs(ref).max()
            assert err < 1e-12, (n, sign, fused, line_fast, err)
            if used:
                took_fused[n] = used
    if fused:
        # 2: fused Rader (prime n > 7 with n-1 = 2^a 3^b 5^c 7^d); 3: composite n with fewer passes
        # through the wide radix set (96 = 6 x 4 x 4, 108 = 6 x 6 x 3, 64 = 8 x 8, ...); 23 - 1 = 2 x 11 stays on plain Rader
        assert {n for n, u in took_fused.items() if u == 2} == {11, 13, 31, 41, 71, 97, 101}
        assert {n for n, u in took_fused.items() if u == 3} == {8, 10, 36, 48, 64, 96, 100, 108, 128}
    else:
        assert not took_fused


def test_thread_counts_and_ragged_batches(emul):
    """Different CTA sizes / line counts hit different work-item splits (and, with real threads,
    would expose a missing barrier as a data race)."""
    rng = np.random.default_rng(5)
    for n in (71, 97, 35):
        for nlines in (1, 2, 9):
            x = rng.standard_normal((nlines, n)) + 1j * rng.standard_normal((nlines, n))
            for nthreads in (1, 3, 16):
                for fused in (0, 1):
                    for line_fast in (0, 1):
                        got, _ = _run(emul, x, -1, fused, line_fast, nthreads)
                        ref = _ref(x, -1)
                        assert abs(got - ref).max() / abs(ref).max() < 1e-12, (n, nlines, nthreads, fused, line_fast)


def test_fused_and_plain_rader_agree_closely(emul):
    """Same arithmetic except the association inside the radix-10 butterfly: agreement ~1e-15."""
    rng = np.random.default_rng(9)
    x = rng.standard_normal((4, 71)) + 1j * rng.standard_normal((4, 71))
    a, _ = _run(emul, x, -1, 0, 1, 8)
    b, used = _run(emul, x, -1, 1, 1, 8)
    assert used == 2
    assert abs(a - b).max() < 1e-13 * abs(a).max()


def test_no_data_races():
    """The same routines under ThreadSanitizer (5 threads per emulated CTA): no race between the
    barrier-separated phases, plain and fused."""
    exe = os.path.join(HERE, 'emul', '_build', 'fft_emul_tsan')
    os.makedirs(os.path.dirname(exe), exist_ok=True)
    cmd = ['g++', '-std=c++20', '-O1', '-g', '-pthread', '-fsanitize=thread', '-w', '-o', exe,
           SRC, os.path.join(HERE, 'emul', 'tsan_main.cpp')]
    built = subprocess.run(cmd, capture_output=True, text=True)
    if built.returncode != 0:
        pytest.skip('g++ -fsanitize=thread is not usable here: ' + built.stderr[-200:])
    run = subprocess.run([exe], capture_output=True, text=True, timeout=600)
    assert run.returncode == 0 and 'tsan run done' in run.stdout, run.stderr[-2000:]
    assert 'ThreadSanitizer' not in run.stderr, run.stderr[-2000:]
