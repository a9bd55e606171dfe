"""ctypes binding of oracle/_ref/libpebbleref.so — the UNMODIFIED reference pebblelib + FFTW 3.3.4
compiled by oracle/build_ref.sh. TEST INFRASTRUCTURE ONLY (checker and CPU baseline)."""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
PATH = os.path.join(_HERE, "_ref", "libpebbleref.so")
_vp, _d, _i = ctypes.c_void_p, ctypes.c_double, ctypes.c_int


def available():
    return os.path.exists(PATH)


_lib = None


def lib():
    global _lib
    if _lib is None:
        L = ctypes.CDLL(PATH)
        sig = {
            "ref_fastfir_create": (_vp, []), "ref_fastfir_destroy": (None, [_vp]),
            "ref_fastfir_setup": (None, [_vp, _d, _d, _d, _d]), "ref_fastfir_process": (_i, [_vp, _i, _vp, _vp]),
            "ref_downconvert_create": (_vp, []), "ref_downconvert_destroy": (None, [_vp]),
            "ref_downconvert_set_data_rate": (_d, [_vp, _d, _d]),
            "ref_downconvert_set_data_rate_simple": (_d, [_vp, _d, _d]),
            "ref_downconvert_set_frequency": (None, [_vp, _d]), "ref_downconvert_set_cw_offset": (None, [_vp, _d]),
            "ref_downconvert_process": (_i, [_vp, _i, _vp, _vp]),
            "ref_fft_create": (_vp, []), "ref_fft_destroy": (None, [_vp]),
            "ref_fft_params": (None, [_vp, ctypes.c_uint, _d, _d, _i, _i]), "ref_fft_size": (_i, [_vp]),
            "ref_fft_forward": (None, [_vp, _vp, _vp, _i]), "ref_fft_inverse": (None, [_vp, _vp, _vp, _i]),
            "ref_fft_spectrum": (_i, [_vp, _vp, _vp, _i]),
            "ref_window": (_d, [_i, _i, _vp]),
            "ref_cfir_create": (_vp, []), "ref_cfir_destroy": (None, [_vp]),
            "ref_cfir_init_lp": (_i, [_vp, _i, _d, _d, _d, _d, _d]), "ref_cfir_init_hp": (_i, [_vp, _i, _d, _d, _d, _d, _d]),
            "ref_cfir_process_cpx": (None, [_vp, _i, _vp, _vp]), "ref_cfir_process_real": (None, [_vp, _i, _vp, _vp]),
            "ref_mixer_create": (_vp, [ctypes.c_uint, ctypes.c_uint]), "ref_mixer_destroy": (None, [_vp]),
            "ref_decimator_create": (_vp, [ctypes.c_uint, ctypes.c_uint]), "ref_decimator_destroy": (None, [_vp]),
            "ref_decimator_build_chain": (ctypes.c_float, [_vp, ctypes.c_uint, ctypes.c_uint, ctypes.c_uint]),
            "ref_decimator_dec_by2_stages": (_i, [_vp]), "ref_decimator_process": (_i, [_vp, _i, _vp, _vp]),
            "ref_mixer_set_frequency": (None, [_vp, _d]), "ref_mixer_process": (_i, [_vp, _i, _vp, _vp]),
            "ref_nco_create": (_vp, [ctypes.c_uint, ctypes.c_uint]), "ref_nco_destroy": (None, [_vp]),
            "ref_nco_set_frequency": (None, [_vp, _d]), "ref_nco_gen_single": (None, [_vp, _vp, ctypes.c_uint, _d, _i]),
            "ref_cpx_op": (None, [_i, _vp, _vp, _vp, _d, _i]),
        }
        for k, (r, a) in sig.items():
            f = getattr(L, k)
            f.restype, f.argtypes = r, a
        _lib = L
    return _lib


class CFastFIR:
    def __init__(self):
        self.h = lib().ref_fastfir_create()

    def SetupParameters(self, lo, hi, off, fs):
        lib().ref_fastfir_setup(self.h, lo, hi, off, fs)

    def ProcessData(self, n, inb, outb):
        return lib().ref_fastfir_process(self.h, n, inb.ctypes.data, outb.ctypes.data)

    def __del__(self):
        lib().ref_fastfir_destroy(self.h)


_lib_long = None


def long_available():
    return os.path.exists(os.path.join(os.path.dirname(PATH), "libpebbleref_long.so"))


def lib_long():
    """The reference CFastFIR compiled with CONV_FFT_SIZE/CONV_FIR_SIZE 131072/65537 (build_ref.sh step 4)."""
    global _lib_long
    if _lib_long is None:
        L = ctypes.CDLL(os.path.join(os.path.dirname(PATH), "libpebbleref_long.so"))
        L.ref_fastfir_create.restype, L.ref_fastfir_create.argtypes = _vp, []
        L.ref_fastfir_destroy.restype, L.ref_fastfir_destroy.argtypes = None, [_vp]
        L.ref_fastfir_setup.restype, L.ref_fastfir_setup.argtypes = None, [_vp, _d, _d, _d, _d]
        L.ref_fastfir_process.restype, L.ref_fastfir_process.argtypes = _i, [_vp, _i, _vp, _vp]
        _lib_long = L
    return _lib_long


class CFastFIRLong:
    """BASELINE config 5: the reference CFastFIR at 131072/65537."""
    FFT_SIZE, FIR_SIZE = 131072, 65537

    def __init__(self):
        self.h = lib_long().ref_fastfir_create()

    def SetupParameters(self, lo, hi, off, fs):
        lib_long().ref_fastfir_setup(self.h, lo, hi, off, fs)

    def ProcessData(self, n, inb, outb):
        return lib_long().ref_fastfir_process(self.h, n, inb.ctypes.data, outb.ctypes.data)

    def __del__(self):
        lib_long().ref_fastfir_destroy(self.h)


class CDownConvert:
    def __init__(self):
        self.h = lib().ref_downconvert_create()

    def SetDataRate(self, r, bw):
        return lib().ref_downconvert_set_data_rate(self.h, r, bw)

    def SetDataRateSimple(self, r, bw):
        return lib().ref_downconvert_set_data_rate_simple(self.h, r, bw)

    def SetFrequency(self, f):
        lib().ref_downconvert_set_frequency(self.h, f)

    def SetCwOffset(self, f):
        lib().ref_downconvert_set_cw_offset(self.h, f)

    def ProcessData(self, n, inb, outb):
        """The reference clobbers its input (downconvert.cpp:309-310): work on a scratch copy."""
        scratch = np.array(inb[:n], dtype=np.complex128, copy=True)
        return lib().ref_downconvert_process(self.h, n, scratch.ctypes.data, outb.ctypes.data)

    def __del__(self):
        lib().ref_downconvert_destroy(self.h)


class FFT:
    def __init__(self):
        self.h = lib().ref_fft_create()

    def fftParams(self, size, dbc, fs, spb, wtype):
        lib().ref_fft_params(self.h, size, dbc, fs, spb, wtype)
        self.size = lib().ref_fft_size(self.h)

    def fftForward(self, inb, outb, n):
        lib().ref_fft_forward(self.h, inb.ctypes.data, outb.ctypes.data, n)

    def fftInverse(self, inb, outb, n):
        lib().ref_fft_inverse(self.h, inb.ctypes.data, outb.ctypes.data, n)

    def fftSpectrum(self, inb, out_db, n):
        return bool(lib().ref_fft_spectrum(self.h, inb.ctypes.data, out_db.ctypes.data, n))

    def __del__(self):
        lib().ref_fft_destroy(self.h)


class CFir:
    def __init__(self):
        self.h = lib().ref_cfir_create()

    def InitLPFilter(self, ntaps, scale, astop, fpass, fstop, fs):
        return lib().ref_cfir_init_lp(self.h, ntaps, scale, astop, fpass, fstop, fs)

    def ProcessFilter(self, n, inb, outb):
        lib().ref_cfir_process_cpx(self.h, n, inb.ctypes.data, outb.ctypes.data)

    def __del__(self):
        lib().ref_cfir_destroy(self.h)


class Mixer:
    def __init__(self, fs, n):
        self.h = lib().ref_mixer_create(int(fs), int(n))
        self.n = n

    def setFrequency(self, f):
        lib().ref_mixer_set_frequency(self.h, f)

    def processBlock(self, inb):
        out = np.empty(self.n, dtype=np.complex128)
        lib().ref_mixer_process(self.h, self.n, inb.ctypes.data, out.ctypes.data)
        return out

    def __del__(self):
        lib().ref_mixer_destroy(self.h)


class Decimator:
    def __init__(self, sampleRate, bufferSize):
        self.h = lib().ref_decimator_create(int(sampleRate), int(bufferSize))

    def buildDecimationChain(self, sampleRateIn, protectBw, sampleRateOut=0):
        return lib().ref_decimator_build_chain(self.h, int(sampleRateIn), int(protectBw), int(sampleRateOut))

    def decBy2Stages(self):
        return lib().ref_decimator_dec_by2_stages(self.h)

    def process(self, inb, outb, n):
        return lib().ref_decimator_process(self.h, n, inb.ctypes.data, outb.ctypes.data)

    def __del__(self):
        lib().ref_decimator_destroy(self.h)


def window(wtype, n):
    w = np.empty(n)
    cg = lib().ref_window(wtype, n, w.ctypes.data)
    return w, cg


class NCO:
    """pebblelib/nco.h: setFrequency (Nyquist safety net, nco.cpp:16-45) + genSingle (nco.cpp:65-81)."""

    def __init__(self, fs, n):
        self.h = lib().ref_nco_create(int(fs), int(n))

    def setFrequency(self, f):
        lib().ref_nco_set_frequency(self.h, float(f))

    def genSingle(self, io, n, gain, mix):
        lib().ref_nco_gen_single(self.h, io.ctypes.data, int(n), float(gain), 1 if mix else 0)

    def __del__(self):
        lib().ref_nco_destroy(self.h)


CPX_OPS = {"copy": 0, "scale": 1, "mult": 2, "add": 3, "mag": 4, "sqrmag": 5, "clear": 6, "decimate": 7}


def cpx_op(name, a, b=None, s=0.0):
    """CpxUtil::{copy,scale,mult,add,mag,sqrMag,clear,decimate}CPX (pebblelib/cpx.cpp:35-112)."""
    a = np.ascontiguousarray(a, dtype=np.complex128)
    n = len(a)
    out = np.zeros((n + int(s) - 1) // int(s) if name == "decimate" else n, dtype=np.complex128)
    b = np.ascontiguousarray(b, dtype=np.complex128) if b is not None else a
    lib().ref_cpx_op(CPX_OPS[name], out.ctypes.data, a.ctypes.data, b.ctypes.data, float(s), n)
    return out
