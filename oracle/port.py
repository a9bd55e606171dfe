"""ctypes binding of oracle/libpebbleoracle.so — the plain-C CPU restatement in oracle/pebble_oracle.c.
TEST INFRASTRUCTURE ONLY (checker; bench.py cpu_baseline kind "port"). Built by pebblesdr_b200/build.py:build_oracle()."""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
PATH = os.path.join(_HERE, "libpebbleoracle.so")
_vp, _d, _i = ctypes.c_void_p, ctypes.c_double, ctypes.c_int
_lib = None


def build():
    src = os.path.join(_HERE, "pebble_oracle.c")
    deps = [src, os.path.join(_HERE, "hb_coefs_c.h"), os.path.join(_HERE, "dec_coefs_c.h")]
    if not os.path.exists(PATH) or os.path.getmtime(PATH) < max(os.path.getmtime(d) for d in deps):
        subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-o", PATH, src, "-lm"])


def lib():
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(PATH)
        sig = {
            "orc_fastfir_create": (_vp, [_i, _i]), "orc_fastfir_destroy": (None, [_vp]),
            "orc_fastfir_setup": (_i, [_vp, _d, _d, _d, _d]), "orc_fastfir_process": (_i, [_vp, _i, _vp, _vp]),
            "orc_downconvert_create": (_vp, []), "orc_downconvert_destroy": (None, [_vp]),
            "orc_downconvert_set_frequency": (None, [_vp, _d]), "orc_downconvert_set_cw_offset": (None, [_vp, _d]),
            "orc_downconvert_set_data_rate": (_d, [_vp, _d, _d, _i]), "orc_downconvert_num_stages": (_i, [_vp]),
            "orc_downconvert_stage_len": (_i, [_vp, _i]), "orc_downconvert_process": (_i, [_vp, _i, _vp, _vp]),
            "orc_window": (_d, [_i, _i, _vp]),
            "orc_fftop_create": (_vp, [_i, _i, _i]), "orc_fftop_destroy": (None, [_vp]),
            "orc_fftop_forward": (None, [_vp, _vp, _vp, _i]), "orc_fftop_inverse": (None, [_vp, _vp, _vp, _i]),
            "orc_fftop_spectrum": (_i, [_vp, _vp, _vp, _i]), "orc_fftop_set_amp_state": (None, [_vp, _vp]),
            "orc_cfir_create": (_vp, []), "orc_cfir_destroy": (None, [_vp]),
            "orc_cfir_init_lp": (_i, [_vp, _i, _d, _d, _d, _d, _d]), "orc_cfir_init_hp": (_i, [_vp, _i, _d, _d, _d, _d, _d]),
            "orc_cfir_taps": (_i, [_vp, _vp]), "orc_cfir_process": (None, [_vp, _i, _vp, _vp]),
            "orc_demod_create": (_vp, [_d]), "orc_demod_destroy": (None, [_vp]),
            "orc_demod_set_am_bandwidth": (None, [_vp, _d]),
            "orc_demod_am": (None, [_vp, _i, _vp, _vp]), "orc_demod_fm2": (None, [_vp, _i, _vp, _vp]),
            "orc_demod_nfm_pll": (None, [_vp, _i, _vp, _vp]),
            "orc_agc_create": (_vp, [_d]), "orc_agc_destroy": (None, [_vp]), "orc_agc_set_mode": (None, [_vp, _i, _i]),
            "orc_agc_manual_gain": (_d, [_vp]), "orc_agc_set_threshold": (None, [_vp, _i]), "orc_agc_process": (None, [_vp, _i, _vp, _vp]),
            "orc_decimator_create": (_vp, [ctypes.c_uint, ctypes.c_uint]), "orc_decimator_destroy": (None, [_vp]),
            "orc_decimator_build_chain": (ctypes.c_float, [_vp, ctypes.c_uint, ctypes.c_uint, ctypes.c_uint]),
            "orc_decimator_dec_by2_stages": (_i, [_vp]), "orc_decimator_num_stages": (_i, [_vp]),
            "orc_decimator_stage": (_i, [_vp, _i, _vp, _vp]), "orc_decimator_process": (_i, [_vp, _i, _vp, _vp]),
            "orc_normalize_iq": (None, [_i, _i, _d, _d, _i, ctypes.c_uint, _vp, _vp]),
            "orc_receiver_create": (_vp, [_d, _i, _d]), "orc_receiver_destroy": (None, [_vp]),
            "orc_receiver_demod_rate": (_d, [_vp]), "orc_receiver_config": (None, [_vp, _d, _d, _d, _i, _i, _i]),
            "orc_receiver_set_am_bandwidth": (None, [_vp, _d]), "orc_receiver_process": (_i, [_vp, _i, _vp, _vp]),
        }
        for k, (r, a) in sig.items():
            f = getattr(L, k)
            f.restype, f.argtypes = r, a
        _lib = L
    return _lib


class _H:
    _dtor = None

    def __del__(self):
        try:
            getattr(lib(), self._dtor)(self.h)
        except Exception:
            pass


class CFastFIR(_H):
    _dtor = "orc_fastfir_destroy"

    def __init__(self, fft_size=2048, fir_size=1025):
        self.h = lib().orc_fastfir_create(fft_size, fir_size)

    def SetupParameters(self, lo, hi, off, fs):
        return lib().orc_fastfir_setup(self.h, lo, hi, off, fs)

    def ProcessData(self, n, inb, outb):
        return lib().orc_fastfir_process(self.h, n, inb.ctypes.data, outb.ctypes.data)


class CDownConvert(_H):
    _dtor = "orc_downconvert_destroy"

    def __init__(self):
        self.h = lib().orc_downconvert_create()

    def SetDataRate(self, r, bw):
        return lib().orc_downconvert_set_data_rate(self.h, r, bw, 0)

    def SetDataRateSimple(self, r, bw):
        return lib().orc_downconvert_set_data_rate(self.h, r, bw, 1)

    def SetFrequency(self, f):
        lib().orc_downconvert_set_frequency(self.h, f)

    def SetCwOffset(self, f):
        lib().orc_downconvert_set_cw_offset(self.h, f)

    def NumStages(self):
        return lib().orc_downconvert_num_stages(self.h)

    def StageLengths(self):
        return [lib().orc_downconvert_stage_len(self.h, i) for i in range(self.NumStages())]

    def ProcessData(self, n, inb, outb):
        scratch = np.array(inb[:n], dtype=np.complex128, copy=True)  # the algorithm clobbers its input
        return lib().orc_downconvert_process(self.h, n, scratch.ctypes.data, outb.ctypes.data)


class FFT(_H):
    _dtor = "orc_fftop_destroy"

    def __init__(self):
        self.h = None

    def fftParams(self, size, dbc, fs, spb, wtype):
        if self.h:
            lib().orc_fftop_destroy(self.h)
        self.h = lib().orc_fftop_create(size, spb, wtype)
        self.size = max(size, 2048)

    def fftForward(self, inb, outb, n):
        lib().orc_fftop_forward(self.h, inb.ctypes.data, outb.ctypes.data, n)

    def fftInverse(self, inb, outb, n):
        lib().orc_fftop_inverse(self.h, inb.ctypes.data, outb.ctypes.data, n)

    def fftSpectrum(self, inb, out_db, n):
        return bool(lib().orc_fftop_spectrum(self.h, inb.ctypes.data, out_db.ctypes.data, n))

    def __del__(self):
        if self.h:
            lib().orc_fftop_destroy(self.h)


class CFir(_H):
    _dtor = "orc_cfir_destroy"

    def __init__(self):
        self.h = lib().orc_cfir_create()

    def InitLPFilter(self, nt, scale, astop, fpass, fstop, fs):
        return lib().orc_cfir_init_lp(self.h, nt, scale, astop, fpass, fstop, fs)

    def InitHPFilter(self, nt, scale, astop, fpass, fstop, fs):
        return lib().orc_cfir_init_hp(self.h, nt, scale, astop, fpass, fstop, fs)

    def GetTaps(self):
        a = np.zeros(75)
        n = lib().orc_cfir_taps(self.h, a.ctypes.data)
        return a[:n].copy()

    def ProcessFilter(self, n, inb, outb):
        lib().orc_cfir_process(self.h, n, inb.ctypes.data, outb.ctypes.data)


class Decimator(_H):
    """Decimator (pebblelib/decimator.cpp), calls of at most 32768-58 samples like the reference."""
    _dtor = "orc_decimator_destroy"

    def __init__(self, sampleRate, bufferSize):
        self.h = lib().orc_decimator_create(int(sampleRate), int(bufferSize))

    def buildDecimationChain(self, sampleRateIn, protectBw, sampleRateOut=0):
        return lib().orc_decimator_build_chain(self.h, int(sampleRateIn), int(protectBw), int(sampleRateOut))

    def decBy2Stages(self):
        return lib().orc_decimator_dec_by2_stages(self.h)

    def chain(self):
        t, d = ctypes.c_int(), ctypes.c_int()
        out = []
        for i in range(lib().orc_decimator_num_stages(self.h)):
            lib().orc_decimator_stage(self.h, i, ctypes.byref(t), ctypes.byref(d))
            out.append((t.value, d.value))
        return out

    def process(self, inb, outb, n):
        return lib().orc_decimator_process(self.h, n, inb.ctypes.data, outb.ctypes.data)


class Demod(_H):
    _dtor = "orc_demod_destroy"

    def __init__(self, rate):
        self.h = lib().orc_demod_create(float(rate))

    def setBandwidth(self, bw):
        lib().orc_demod_set_am_bandwidth(self.h, bw)

    def _run(self, fn, inb, n):
        out = np.empty(n, dtype=np.complex128)
        getattr(lib(), fn)(self.h, n, inb.ctypes.data, out.ctypes.data)
        return out

    def am(self, inb, n):
        return self._run("orc_demod_am", inb, n)

    def fm2(self, inb, n):
        return self._run("orc_demod_fm2", inb, n)

    def nfm_pll(self, inb, n):
        return self._run("orc_demod_nfm_pll", inb, n)


class AGC(_H):
    _dtor = "orc_agc_destroy"

    def __init__(self, rate):
        self.h = lib().orc_agc_create(float(rate))

    def setAgcMode(self, m, thr):
        lib().orc_agc_set_mode(self.h, m, thr)

    def setAgcThreshold(self, g):
        lib().orc_agc_set_threshold(self.h, g)

    def processBlock(self, inb, n):
        out = np.empty(n, dtype=np.complex128)
        lib().orc_agc_process(self.h, n, inb.ctypes.data, out.ctypes.data)
        return out


class Receiver(_H):
    _dtor = "orc_receiver_destroy"

    def __init__(self, rate, fpb, max_bw):
        self.h = lib().orc_receiver_create(float(rate), int(fpb), float(max_bw))
        self.demod_rate = lib().orc_receiver_demod_rate(self.h)
        self.fpb = fpb

    def config(self, mixer, lo, hi, demod_mode, agc_mode=0, agc_threshold=20):
        lib().orc_receiver_config(self.h, mixer, lo, hi, demod_mode, agc_mode, agc_threshold)

    def setAmBandwidth(self, bw):
        lib().orc_receiver_set_am_bandwidth(self.h, bw)

    def processIQData(self, inb, n):
        """n <= 65536 per call. Returns the demodulated samples produced by this call."""
        scratch = np.array(inb[:n], dtype=np.complex128, copy=True)
        out = np.empty(n + 2 * self.fpb, dtype=np.complex128)
        k = lib().orc_receiver_process(self.h, n, scratch.ctypes.data, out.ctypes.data)
        return out[:k].copy()


def window(wtype, n):
    w = np.zeros(n)
    cg = lib().orc_window(wtype, n, w.ctypes.data)
    return w, cg


FMT_CPX8, FMT_CPXU8, FMT_CPX16, FMT_CPXFLOAT, FMT_CPX = range(5)
_FMT_DTYPE = {FMT_CPX8: np.int8, FMT_CPXU8: np.uint8, FMT_CPX16: np.int16, FMT_CPXFLOAT: np.float32, FMT_CPX: np.float64}


def normalizeIQ(fmt, raw, order=0, user_gain=1.0, norm_gain=1.0, reverse=False):
    """DeviceInterfaceBase::normalizeIQ block forms (deviceinterfacebase.cpp:648-878). raw: interleaved I,Q array."""
    raw = np.ascontiguousarray(raw, dtype=_FMT_DTYPE[fmt]).reshape(-1)
    n = raw.size // 2
    out = np.empty(n, dtype=np.complex128)
    lib().orc_normalize_iq(fmt, order, user_gain, norm_gain, 1 if reverse else 0, n, raw.ctypes.data, out.ctypes.data)
    return out
