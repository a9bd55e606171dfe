"""ctypes binding of oracle/_ref/libpebbleref_app.so (the UNMODIFIED application-side steps of the reference: AGC,
Demod_AM/NFM/SAM/WFM, SignalSpectrum, SignalStrength, DCRemoval, IQBalance, NoiseBlanker, CFractResampler, CIir,
normalizeIQ — oracle/build_ref.sh step 5), of libpebbleref_dec12.so (CDownConvert with MAX_DECSTAGES 12, step 6) and of
libpebbleref_cute.so (the Ooura FFT backend, step 7). TEST INFRASTRUCTURE ONLY (checker and CPU baseline)."""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_DIR = os.path.join(_HERE, "_ref")
PATH = os.path.join(_DIR, "libpebbleref_app.so")
PATH_DEC12 = os.path.join(_DIR, "libpebbleref_dec12.so")
PATH_CUTE = os.path.join(_DIR, "libpebbleref_cute.so")
_vp, _d, _i, _u, _f, _ll = ctypes.c_void_p, ctypes.c_double, ctypes.c_int, ctypes.c_uint, ctypes.c_float, ctypes.c_longlong


def available():
    return os.path.exists(PATH)


def dec12_available():
    return os.path.exists(PATH_DEC12)


def cute_available():
    return os.path.exists(PATH_CUTE)


_lib = None


def lib():
    global _lib
    if _lib is None:
        L = ctypes.CDLL(PATH)
        sig = {
            "ref_set_clock_ms": (None, [_ll]),
            "ref_agc_create": (_vp, [_u, _u]), "ref_agc_destroy": (None, [_vp]), "ref_agc_set_mode": (None, [_vp, _i, _i]),
            "ref_agc_set_threshold": (None, [_vp, _i]), "ref_agc_get_threshold": (_i, [_vp]),
            "ref_agc_process": (None, [_vp, _i, _vp, _vp]),
            "ref_demod_am_create": (_vp, [_i, _i]), "ref_demod_am_destroy": (None, [_vp]),
            "ref_demod_am_set_bandwidth": (None, [_vp, _d]), "ref_demod_am_process_filtered": (None, [_vp, _i, _vp, _vp]),
            "ref_demod_am_process": (None, [_vp, _i, _vp, _vp]),
            "ref_demod_nfm_create": (_vp, [_i, _i]), "ref_demod_nfm_destroy": (None, [_vp]),
            "ref_demod_nfm_process_nco": (None, [_vp, _i, _vp, _vp]), "ref_demod_nfm_process_fm2": (None, [_vp, _i, _vp, _vp]),
            "ref_demod_sam_create": (_vp, [_i, _i]), "ref_demod_sam_destroy": (None, [_vp]),
            "ref_demod_sam_process": (None, [_vp, _i, _vp, _vp]),
            "ref_demod_wfm_create": (_vp, [_i, _i]), "ref_demod_wfm_destroy": (None, [_vp]),
            "ref_demod_wfm_set_sample_rate": (_d, [_vp, _d, _d, _i]),
            "ref_demod_wfm_process_mono": (_i, [_vp, _i, _vp, _vp]), "ref_demod_wfm_process_stereo": (_i, [_vp, _i, _vp, _vp]),
            "ref_demod_wfm_stereo_lock": (_i, [_vp, _vp]), "ref_demod_wfm_next_rds_group": (_i, [_vp, _vp]),
            "ref_demod_wfm_rds_tap": (_i, [_vp, _i, _vp]), "ref_demod_wfm_rds_rate": (_d, [_vp]),
            "ref_demod_wfm_push_bits": (None, [_vp, _vp, _i]),
            "ref_spectrum_create": (_vp, [_u, _u, _u, _i, _i, _i]), "ref_spectrum_destroy": (None, [_vp]),
            "ref_spectrum_set_hires": (None, [_vp, _i]), "ref_spectrum_set_updates_per_sec": (None, [_vp, _i]),
            "ref_spectrum_set_sample_rate": (None, [_vp, _u, _u]),
            "ref_spectrum_unprocessed": (None, [_vp, _vp, _i]), "ref_spectrum_zoomed": (None, [_vp, _vp, _i]),
            "ref_spectrum_bin_count": (_i, [_vp]), "ref_spectrum_overload": (_i, [_vp]),
            "ref_spectrum_update_pending": (_i, [_vp]), "ref_spectrum_ack_update": (None, [_vp]),
            "ref_spectrum_overrun": (_i, [_vp]), "ref_spectrum_get_unprocessed": (None, [_vp, _vp]),
            "ref_sigstrength_create": (_vp, [_u, _u]), "ref_sigstrength_destroy": (None, [_vp]),
            "ref_sigstrength_fd_estimate": (_d, [_vp, _vp, _i, _u, _f, _f, _d]), "ref_sigstrength_values": (None, [_vp, _vp]),
            "ref_dcremoval_create": (_vp, [_u, _u]), "ref_dcremoval_destroy": (None, [_vp]),
            "ref_dcremoval_enable": (None, [_vp, _i]), "ref_dcremoval_process": (None, [_vp, _i, _vp, _vp]),
            "ref_iqbalance_create": (_vp, [_u, _u]), "ref_iqbalance_destroy": (None, [_vp]),
            "ref_iqbalance_config": (None, [_vp, _i, _d, _d]), "ref_iqbalance_process": (_i, [_vp, _i, _vp, _vp]),
            "ref_noiseblanker_create": (_vp, [_u, _u]), "ref_noiseblanker_destroy": (None, [_vp]),
            "ref_noiseblanker_enable": (None, [_vp, _i, _i]), "ref_noiseblanker_process": (_i, [_vp, _i, _vp, _vp]),
            "ref_noiseblanker_process2": (_i, [_vp, _i, _vp, _vp]),
            "ref_resampler_create": (_vp, [_i]), "ref_resampler_destroy": (None, [_vp]),
            "ref_resampler_resample_cpx": (_i, [_vp, _i, _d, _vp, _vp]),
            "ref_resampler_resample_stereo16": (_i, [_vp, _i, _d, _vp, _vp, _d]),
            "ref_iir_create": (_vp, []), "ref_iir_destroy": (None, [_vp]), "ref_iir_init": (None, [_vp, _i, _d, _d, _d]),
            "ref_iir_process_cpx": (None, [_vp, _i, _vp, _vp]), "ref_iir_process_real": (None, [_vp, _i, _vp, _vp]),
            "ref_normalize_iq": (None, [_i, _i, _d, _d, _i, _u, _vp, _vp]),
        }
        for k, (r, a) in sig.items():
            f = getattr(L, k)
            f.restype, f.argtypes = r, a
        _lib = L
    return _lib


def set_clock_ms(t):
    """The QElapsedTimer stand-in reads this fake clock: tests drive the reference's update gates with it."""
    lib().ref_set_clock_ms(int(t))


class _H:
    _dtor = None

    def __del__(self):
        try:
            getattr(lib(), self._dtor)(self.h)
        except Exception:
            pass


def _c(x):
    return np.ascontiguousarray(x, dtype=np.complex128)


class AGC(_H):
    """application/agc.h. processBlock handles exactly bufferSize samples per call (agc.cpp:84-207)."""
    _dtor = "ref_agc_destroy"
    AGC_OFF, ACG_FAST, AGC_MED, AGC_SLOW, AGC_LONG = range(5)

    def __init__(self, sampleRate, bufferSize):
        self.n = int(bufferSize)
        self.h = lib().ref_agc_create(int(sampleRate), self.n)

    def setAgcMode(self, m, thr):
        lib().ref_agc_set_mode(self.h, m, thr)

    def setAgcThreshold(self, g):
        lib().ref_agc_set_threshold(self.h, g)

    def processBlock(self, inb):
        inb = _c(inb)
        assert len(inb) == self.n
        out = np.empty(self.n, dtype=np.complex128)
        lib().ref_agc_process(self.h, self.n, inb.ctypes.data, out.ctypes.data)
        return out

    def run(self, x):
        """Whole stream in bufferSize blocks (len(x) must be a multiple)."""
        x = _c(x)
        assert len(x) % self.n == 0
        return np.concatenate([self.processBlock(x[i:i + self.n]) for i in range(0, len(x), self.n)])


class _Demod(_H):
    def _run(self, fn, inb, n=None):
        inb = _c(inb)
        n = len(inb) if n is None else n
        out = np.empty(n, dtype=np.complex128)
        getattr(lib(), fn)(self.h, n, inb.ctypes.data, out.ctypes.data)
        return out


class Demod_AM(_Demod):
    _dtor = "ref_demod_am_destroy"

    def __init__(self, rate, n):
        self.h = lib().ref_demod_am_create(int(rate), int(n))

    def setBandwidth(self, bw):
        lib().ref_demod_am_set_bandwidth(self.h, float(bw))

    def processBlockFiltered(self, inb):
        return self._run("ref_demod_am_process_filtered", inb)

    def processBlock(self, inb):
        return self._run("ref_demod_am_process", inb)


class Demod_NFM(_Demod):
    _dtor = "ref_demod_nfm_destroy"

    def __init__(self, rate, n):
        self.h = lib().ref_demod_nfm_create(int(rate), int(n))

    def processBlockNCO(self, inb):
        return self._run("ref_demod_nfm_process_nco", inb)

    def processBlockFM2(self, inb):
        """NOTE: the last sample is a function-static shared by every object in the process (demod_nfm.cpp:129)."""
        return self._run("ref_demod_nfm_process_fm2", inb)


class Demod_SAM(_Demod):
    _dtor = "ref_demod_sam_destroy"

    def __init__(self, rate, n):
        self.h = lib().ref_demod_sam_create(int(rate), int(n))

    def processBlock(self, inb):
        return self._run("ref_demod_sam_process", inb)


class Demod_WFM(_H):
    _dtor = "ref_demod_wfm_destroy"

    def __init__(self, rate, n):
        self.h = lib().ref_demod_wfm_create(int(rate), int(n))

    def setSampleRate(self, rate, out_rate=0.0, us=True):
        return lib().ref_demod_wfm_set_sample_rate(self.h, float(rate), float(out_rate), 1 if us else 0)

    def _run(self, fn, inb):
        inb = _c(inb)
        out = np.empty(len(inb), dtype=np.complex128)
        k = getattr(lib(), fn)(self.h, len(inb), inb.ctypes.data, out.ctypes.data)
        return out[:k].copy()

    def processDataMono(self, inb):
        return self._run("ref_demod_wfm_process_mono", inb)

    def processDataStereo(self, inb):
        return self._run("ref_demod_wfm_process_stereo", inb)

    def getStereoLock(self):
        v = ctypes.c_int(0)
        ch = lib().ref_demod_wfm_stereo_lock(self.h, ctypes.byref(v))
        return bool(ch), v.value

    def getNextRdsGroupData(self):
        b = (ctypes.c_ushort * 4)()
        if lib().ref_demod_wfm_next_rds_group(self.h, b):
            return tuple(b)
        return None

    def pushRdsBits(self, bits):
        """Demod_WFM::processNewRdsBit for every element of `bits` (data bits, after the differential decode)."""
        b = np.ascontiguousarray(bits, dtype=np.uint8)
        lib().ref_demod_wfm_push_bits(self.h, b.ctypes.data, len(b))

    def rdsRate(self):
        return lib().ref_demod_wfm_rds_rate(self.h)

    def rdsTap(self, length):
        """m_RdsRaw after the last processDataStereo call (the debug pairs the reference leaves there, demod_wfm.cpp:336-352)."""
        out = np.empty(int(length), dtype=np.complex128)
        lib().ref_demod_wfm_rds_tap(self.h, int(length), out.ctypes.data)
        return out


class SignalSpectrum(_H):
    """application/signalspectrum.h with `global->settings` = (bins, hires_bins, updates_per_sec)."""
    _dtor = "ref_spectrum_destroy"

    def __init__(self, sampleRate, hiResSampleRate, bufferSize, bins=4096, hires_bins=2048, updates_per_sec=10):
        self.h = lib().ref_spectrum_create(int(sampleRate), int(hiResSampleRate), int(bufferSize), bins, hires_bins, updates_per_sec)
        self.bins = bins

    def setHiRes(self, on):
        lib().ref_spectrum_set_hires(self.h, 1 if on else 0)

    def setUpdatesPerSec(self, u):
        lib().ref_spectrum_set_updates_per_sec(self.h, int(u))

    def setSampleRate(self, fs, hires_fs):
        lib().ref_spectrum_set_sample_rate(self.h, int(fs), int(hires_fs))

    def unprocessed(self, inb, n):
        inb = _c(inb)
        lib().ref_spectrum_unprocessed(self.h, inb.ctypes.data, n)

    def zoomed(self, inb, n):
        inb = _c(inb)
        lib().ref_spectrum_zoomed(self.h, inb.ctypes.data, n)

    def getOverload(self):
        return bool(lib().ref_spectrum_overload(self.h))

    def updatePending(self):
        return bool(lib().ref_spectrum_update_pending(self.h))

    def ackUpdate(self):
        lib().ref_spectrum_ack_update(self.h)

    def overrun(self):
        return lib().ref_spectrum_overrun(self.h)

    def getUnprocessed(self):
        out = np.empty(self.bins)
        lib().ref_spectrum_get_unprocessed(self.h, out.ctypes.data)
        return out


class SignalStrength(_H):
    _dtor = "ref_sigstrength_destroy"

    def __init__(self, sampleRate, bufferSize):
        self.h = lib().ref_sigstrength_create(int(sampleRate), int(bufferSize))

    def fdEstimate(self, spectrum, bins, fs, lo, hi, mixer):
        s = np.ascontiguousarray(spectrum, dtype=np.float64)
        return lib().ref_sigstrength_fd_estimate(self.h, s.ctypes.data, int(bins), int(fs), float(lo), float(hi), float(mixer))

    def values(self):
        """(peakDb, avgDb, snrDb, floorDb, extValue)"""
        v = np.empty(5)
        lib().ref_sigstrength_values(self.h, v.ctypes.data)
        return v


class _Step(_H):
    def _proc(self, fn, inb):
        inb = _c(inb)
        out = np.empty(len(inb), dtype=np.complex128)
        r = getattr(lib(), fn)(self.h, len(inb), inb.ctypes.data, out.ctypes.data)
        return out, r


class DCRemoval(_Step):
    _dtor = "ref_dcremoval_destroy"

    def __init__(self, fs, n):
        self.h = lib().ref_dcremoval_create(int(fs), int(n))

    def enableStep(self, on):
        lib().ref_dcremoval_enable(self.h, 1 if on else 0)

    def process(self, inb):
        return self._proc("ref_dcremoval_process", inb)[0]


class IQBalance(_Step):
    """ProcessBlock handles exactly bufferSize samples (iqbalance.cpp:67-88)."""
    _dtor = "ref_iqbalance_destroy"

    def __init__(self, fs, n):
        self.h = lib().ref_iqbalance_create(int(fs), int(n))

    def config(self, enabled, gain=1.0, phase=0.0):
        lib().ref_iqbalance_config(self.h, 1 if enabled else 0, gain, phase)

    def ProcessBlock(self, inb):
        return self._proc("ref_iqbalance_process", inb)


class NoiseBlanker(_Step):
    _dtor = "ref_noiseblanker_destroy"

    def __init__(self, fs, n):
        self.h = lib().ref_noiseblanker_create(int(fs), int(n))

    def enable(self, nb, nb2):
        lib().ref_noiseblanker_enable(self.h, 1 if nb else 0, 1 if nb2 else 0)

    def ProcessBlock(self, inb):
        return self._proc("ref_noiseblanker_process", inb)

    def ProcessBlock2(self, inb):
        return self._proc("ref_noiseblanker_process2", inb)


class CFractResampler(_H):
    _dtor = "ref_resampler_destroy"

    def __init__(self, maxInputSize):
        self.h = lib().ref_resampler_create(int(maxInputSize))

    def Resample(self, inb, rate):
        inb = _c(inb)
        out = np.empty(int(len(inb) / rate) + 8, dtype=np.complex128)
        k = lib().ref_resampler_resample_cpx(self.h, len(inb), float(rate), inb.ctypes.data, out.ctypes.data)
        return out[:k].copy()

    def ResampleStereo16(self, inb, rate, gain):
        inb = _c(inb)
        out = np.empty((int(len(inb) / rate) + 8, 2), dtype=np.int16)
        k = lib().ref_resampler_resample_stereo16(self.h, len(inb), float(rate), inb.ctypes.data, out.ctypes.data, float(gain))
        return out[:k].copy()


class CIir(_H):
    _dtor = "ref_iir_destroy"
    LP, HP, BP, BR = range(4)

    def __init__(self, kind, f0, q, fs):
        self.h = lib().ref_iir_create()
        lib().ref_iir_init(self.h, kind, float(f0), float(q), float(fs))

    def ProcessFilter(self, inb):
        if np.iscomplexobj(inb):
            inb = _c(inb)
            out = np.empty(len(inb), dtype=np.complex128)
            lib().ref_iir_process_cpx(self.h, len(inb), inb.ctypes.data, out.ctypes.data)
        else:
            inb = np.ascontiguousarray(inb, dtype=np.float64)
            out = np.empty(len(inb))
            lib().ref_iir_process_real(self.h, len(inb), inb.ctypes.data, out.ctypes.data)
        return out


FMT_CPX8, FMT_CPXU8, FMT_CPX16, FMT_CPXFLOAT, FMT_CPX = range(5)
_FMT_DTYPE = {FMT_CPX8: np.int8, FMT_CPXU8: np.uint8, FMT_CPX16: np.int16, FMT_CPXFLOAT: np.float32, FMT_CPX: np.float64}


def normalizeIQ(fmt, raw, order=0, user_gain=1.0, norm_gain=1.0, reverse=False):
    """DeviceInterfaceBase::normalizeIQ block forms (deviceinterfacebase.cpp:648-878). raw: interleaved I,Q array."""
    raw = np.ascontiguousarray(raw, dtype=_FMT_DTYPE[fmt]).reshape(-1)
    n = raw.size // 2
    out = np.empty(n, dtype=np.complex128)
    lib().ref_normalize_iq(fmt, order, user_gain, norm_gain, 1 if reverse else 0, n, raw.ctypes.data, out.ctypes.data)
    return out


# ---- MAX_DECSTAGES 12 build of CDownConvert (BASELINE config 4) --------------------------------------------------------
_lib12 = None


def lib_dec12():
    global _lib12
    if _lib12 is None:
        L = ctypes.CDLL(PATH_DEC12)
        for k, (r, a) in {
            "ref_downconvert_create": (_vp, []), "ref_downconvert_destroy": (None, [_vp]),
            "ref_downconvert_set_data_rate": (_d, [_vp, _d, _d]), "ref_downconvert_set_frequency": (None, [_vp, _d]),
            "ref_downconvert_process": (_i, [_vp, _i, _vp, _vp]),
        }.items():
            f = getattr(L, k)
            f.restype, f.argtypes = r, a
        _lib12 = L
    return _lib12


class CDownConvert12:
    """The reference CDownConvert compiled with MAX_DECSTAGES 12 (downconvert.h:23): runs the 10-stage 61.44 MSPS chain."""

    def __init__(self):
        self.h = lib_dec12().ref_downconvert_create()

    def SetDataRate(self, r, bw):
        return lib_dec12().ref_downconvert_set_data_rate(self.h, r, bw)

    def SetFrequency(self, f):
        lib_dec12().ref_downconvert_set_frequency(self.h, f)

    def ProcessData(self, n, inb, outb):
        scratch = np.array(inb[:n], dtype=np.complex128, copy=True)   # the reference clobbers its input
        return lib_dec12().ref_downconvert_process(self.h, n, scratch.ctypes.data, outb.ctypes.data)

    def __del__(self):
        try:
            lib_dec12().ref_downconvert_destroy(self.h)
        except Exception:
            pass


# ---- Ooura backend (class CFft, -DUSE_FFTCUTE) ----------------------------------------------------------------------------
_libcute = None


def lib_cute():
    global _libcute
    if _libcute is None:
        L = ctypes.CDLL(PATH_CUTE)
        for k, (r, a) in {
            "ref_fft_create": (_vp, []), "ref_fft_destroy": (None, [_vp]),
            "ref_fft_params": (None, [_vp, _u, _d, _d, _i, _i]), "ref_fft_size": (_i, [_vp]),
            "ref_fft_forward": (None, [_vp, _vp, _vp, _i]), "ref_fft_inverse": (None, [_vp, _vp, _vp, _i]),
            "ref_fft_spectrum": (_i, [_vp, _vp, _vp, _i]),
        }.items():
            f = getattr(L, k)
            f.restype, f.argtypes = r, a
        _libcute = L
    return _libcute


class CFft:
    """FFT::factory() of a -DUSE_FFTCUTE build = class CFft (pebblelib/fftcute.cpp)."""

    def __init__(self):
        self.h = lib_cute().ref_fft_create()

    def fftParams(self, size, dbc, fs, spb, wtype):
        lib_cute().ref_fft_params(self.h, size, dbc, fs, spb, wtype)
        self.size = lib_cute().ref_fft_size(self.h)

    def fftForward(self, inb, outb, n):
        lib_cute().ref_fft_forward(self.h, inb.ctypes.data, outb.ctypes.data, n)

    def fftInverse(self, inb, outb, n):
        lib_cute().ref_fft_inverse(self.h, inb.ctypes.data, outb.ctypes.data, n)

    def fftSpectrum(self, inb, out_db, n):
        return bool(lib_cute().ref_fft_spectrum(self.h, inb.ctypes.data, out_db.ctypes.data, n))

    def __del__(self):
        try:
            lib_cute().ref_fft_destroy(self.h)
        except Exception:
            pass
