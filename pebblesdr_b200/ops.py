"""Host-side mirror of the reference operator interface for the hot path, over the C ABI.

Each class keeps the reference's method names and argument meaning:
  CFastFIR      pebblelib/fastfir.h:50-77     SetupParameters / ProcessData
  CDownConvert  pebblelib/downconvert.h:28-37 SetDataRate / SetFrequency / SetCwOffset / ProcessData
`ProcessData(n, in, out)` takes numpy complex128 arrays (host, the reference call shape) or torch
complex128 CUDA tensors (device-resident throughput path) and returns the output sample count."""
import ctypes

import numpy as np

from ._lib import lib, check, PebbleError

try:  # torch is only needed for device-resident buffers
    import torch
except Exception:  # pragma: no cover
    torch = None


def _is_cuda_tensor(x):
    return torch is not None and isinstance(x, torch.Tensor) and x.is_cuda


def _host_ptr(a, n, writable=False):
    if not isinstance(a, np.ndarray) or a.dtype != np.complex128 or not a.flags.c_contiguous or a.size < n:
        raise PebbleError("expected a C-contiguous numpy complex128 array with >= %d samples" % n)
    if writable and not a.flags.writeable:
        raise PebbleError("output array is read-only")
    return a.ctypes.data


def _dev_ptr(t, n):
    if t.dtype != torch.complex128 or not t.is_contiguous() or t.numel() < n:
        raise PebbleError("expected a contiguous torch complex128 CUDA tensor with >= %d samples" % n)
    return t.data_ptr()


def _stream_ptr():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


# Device sample formats (include/pebble_cuda.h PEBBLE_IQ_*; pebblelib/cpx.h:43-96): raw buffers hold interleaved I,Q
IQ_CPX, IQ_CS16, IQ_CU8, IQ_CS8, IQ_CF32 = range(5)
IQO_IQ, IQO_QI, IQO_IONLY, IQO_QONLY = range(4)
_IQ_NP = {IQ_CPX: np.complex128, IQ_CS16: np.int16, IQ_CU8: np.uint8, IQ_CS8: np.int8, IQ_CF32: np.float32}


def _raw_ptr(x, n, fmt):
    """Pointer to n raw samples of format fmt: numpy array or torch tensor (CPU pinned or CUDA) of the matching
    element type, 2 elements per sample (1 for complex128)."""
    if fmt == IQ_CPX:
        return _dev_ptr(x, n) if (torch is not None and isinstance(x, torch.Tensor)) else _host_ptr(x, n)
    need = 2 * n
    if torch is not None and isinstance(x, torch.Tensor):
        tdt = {IQ_CS16: torch.int16, IQ_CU8: torch.uint8, IQ_CS8: torch.int8, IQ_CF32: torch.float32}[fmt]
        if x.dtype != tdt or not x.is_contiguous() or x.numel() < need:
            raise PebbleError("expected a contiguous torch %s tensor with >= %d elements" % (tdt, need))
        return x.data_ptr()
    if not isinstance(x, np.ndarray) or x.dtype != _IQ_NP[fmt] or not x.flags.c_contiguous or x.size < need:
        raise PebbleError("expected a C-contiguous numpy %s array with >= %d elements" % (_IQ_NP[fmt].__name__, need))
    return x.ctypes.data


def normalizeIQ(fmt, raw, order=IQO_IQ, userIQGain=1.0, normalizeIQGain=1.0, reverse=False, out=None):
    """DeviceInterfaceBase::normalizeIQ block forms (pebblelib/deviceinterfacebase.cpp:648-878) on the GPU.
    raw: interleaved I,Q numpy array (host) or torch CUDA tensor (device); returns complex128 of the same kind."""
    n = (raw.numel() if _is_cuda_tensor(raw) else raw.size) // (1 if fmt == IQ_CPX else 2)
    if _is_cuda_tensor(raw):
        if out is None:
            out = torch.empty(n, dtype=torch.complex128, device=raw.device)
        check(lib.pebble_normalize_iq_device(fmt, order, userIQGain, normalizeIQGain, int(bool(reverse)), n,
                                             _raw_ptr(raw, n, fmt), _dev_ptr(out, n), _stream_ptr()), "normalizeIQ")
        return out
    if out is None:
        out = np.empty(n, dtype=np.complex128)
    check(lib.pebble_normalize_iq(fmt, order, userIQGain, normalizeIQGain, int(bool(reverse)), n, _raw_ptr(raw, n, fmt),
                                  _host_ptr(out, n, True)), "normalizeIQ")
    return out


class _Handle:
    _destroy = None

    def __init__(self, h, what):
        if not h:
            raise PebbleError("%s: create failed: %s" % (what, lib.pebble_last_error().decode()))
        self._h = ctypes.c_void_p(h)

    def close(self):
        if getattr(self, "_h", None):
            type(self)._destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class CFastFIR(_Handle):
    """Overlap-save FFT band-pass (pebblelib/fastfir.cpp:77-334)."""
    _destroy = staticmethod(lib.pebble_fastfir_destroy)

    def __init__(self, fft_size=2048, fir_size=1025):
        super().__init__(lib.pebble_fastfir_create_sized(fft_size, fir_size), "CFastFIR")
        self.fft_size, self.fir_size = fft_size, fir_size

    def SetupParameters(self, FLoCut, FHiCut, Offset, SampleRate):
        return check(lib.pebble_fastfir_setup(self._h, FLoCut, FHiCut, Offset, SampleRate), "SetupParameters")

    def ProcessData(self, InLength, InBuf, OutBuf):
        if _is_cuda_tensor(InBuf):
            return check(lib.pebble_fastfir_process_device(self._h, InLength, _dev_ptr(InBuf, InLength),
                                                           _dev_ptr(OutBuf, self.OutCount(InLength)), _stream_ptr()),
                         "ProcessData")
        return check(lib.pebble_fastfir_process(self._h, InLength, _host_ptr(InBuf, InLength),
                                                _host_ptr(OutBuf, self.OutCount(InLength), True)), "ProcessData")

    def OutCount(self, InLength):
        return check(lib.pebble_fastfir_out_count(self._h, InLength), "OutCount")

    def SetFusedScale(self, in_scale, out_scale):
        check(lib.pebble_fastfir_set_fused_scale(self._h, in_scale, out_scale))

    def GetFilter(self):
        out = np.empty(self.fft_size, dtype=np.complex128)
        check(lib.pebble_fastfir_get_filter(self._h, out.ctypes.data), "GetFilter")
        return out

    def GetState(self):
        n = check(lib.pebble_fastfir_state_size(self._h))
        buf = np.empty(n, dtype=np.uint8)
        check(lib.pebble_fastfir_get_state(self._h, buf.ctypes.data, n), "GetState")
        return buf

    def SetState(self, buf):
        check(lib.pebble_fastfir_set_state(self._h, buf.ctypes.data, buf.size), "SetState")


class CDownConvert(_Handle):
    """NCO mix + decimate-by-2 cascade (pebblelib/downconvert.cpp:64-535)."""
    _destroy = staticmethod(lib.pebble_downconvert_destroy)

    def __init__(self):
        super().__init__(lib.pebble_downconvert_create(), "CDownConvert")
        self._fmt = IQ_CPX

    def SetInputFormat(self, fmt, order=IQO_IQ, userIQGain=1.0, normalizeIQGain=1.0):
        """ProcessData then takes raw device samples of this format (normalizeIQ fused into the mixer load)."""
        check(lib.pebble_downconvert_set_input_format(self._h, fmt, order, userIQGain, normalizeIQGain), "SetInputFormat")
        self._fmt = fmt

    def SetDataRate(self, InRate, MaxBW):
        return lib.pebble_downconvert_set_data_rate(self._h, InRate, MaxBW)

    def SetDataRateSimple(self, InRate, MaxBW):
        return lib.pebble_downconvert_set_data_rate_simple(self._h, InRate, MaxBW)

    def SetFrequency(self, NcoFreq):
        lib.pebble_downconvert_set_frequency(self._h, NcoFreq)

    def SetCwOffset(self, offset):
        lib.pebble_downconvert_set_cw_offset(self._h, offset)

    def NumStages(self):
        return check(lib.pebble_downconvert_num_stages(self._h))

    def StageLengths(self):
        a = (ctypes.c_int * 16)()
        n = check(lib.pebble_downconvert_stage_lengths(self._h, a, 16))
        return list(a[:n])

    def ProcessData(self, InLength, pInData, pOutData):
        nout = InLength >> self.NumStages()
        if _is_cuda_tensor(pInData):
            return check(lib.pebble_downconvert_process_device(self._h, InLength, _raw_ptr(pInData, InLength, self._fmt),
                                                               _dev_ptr(pOutData, nout), _stream_ptr()), "ProcessData")
        return check(lib.pebble_downconvert_process(self._h, InLength, _raw_ptr(pInData, InLength, self._fmt),
                                                    _host_ptr(pOutData, nout, True)), "ProcessData")

    def GetState(self):
        n = check(lib.pebble_downconvert_state_size(self._h))
        buf = np.empty(n, dtype=np.uint8)
        check(lib.pebble_downconvert_get_state(self._h, buf.ctypes.data, n), "GetState")
        return buf

    def SetState(self, buf):
        check(lib.pebble_downconvert_set_state(self._h, buf.ctypes.data, buf.size), "SetState")


def decimator_plan(sampleRateIn, protectBw, sampleRateOut=0):
    """Host-only Decimator::buildDecimationChain (decimator.cpp:64-149): ([(taps, decimate)], rate, decBy2Stages);
    taps == 0 is the CIC3 stage."""
    t, d = (ctypes.c_int * 16)(), (ctypes.c_int * 16)()
    r, s2 = ctypes.c_float(), ctypes.c_int()
    n = lib.pebble_decimator_plan(int(sampleRateIn), int(protectBw), int(sampleRateOut), t, d, 16, ctypes.byref(r),
                                  ctypes.byref(s2))
    return [(t[i], d[i]) for i in range(n)], r.value, s2.value


class Decimator(_Handle):
    """The alternative decimation chain (pebblelib/decimator.h:228-284, decimator.cpp:6-1339)."""
    _destroy = staticmethod(lib.pebble_decimator_destroy)

    def __init__(self, _sampleRate, _bufferSize):
        super().__init__(lib.pebble_decimator_create(int(_sampleRate), int(_bufferSize)), "Decimator")

    def buildDecimationChain(self, _sampleRateIn, _protectBw, _sampleRateOut=0):
        return lib.pebble_decimator_build_chain(self._h, int(_sampleRateIn), int(_protectBw), int(_sampleRateOut))

    def decBy2Stages(self):
        return check(lib.pebble_decimator_dec_by2_stages(self._h))

    def totalDecimation(self):
        return check(lib.pebble_decimator_total_decimation(self._h))

    def chain(self):
        t, d = (ctypes.c_int * 16)(), (ctypes.c_int * 16)()
        n = check(lib.pebble_decimator_chain(self._h, t, d, 16))
        return [(t[i], d[i]) for i in range(n)]

    def outCount(self, n):
        return lib.pebble_decimator_out_count(self._h, n)

    def process(self, _in, _out, _numSamples):
        nout = check(lib.pebble_decimator_out_count(self._h, _numSamples), "process")
        if _is_cuda_tensor(_in):
            return check(lib.pebble_decimator_process_device(self._h, _numSamples, _dev_ptr(_in, _numSamples),
                                                             _dev_ptr(_out, nout), _stream_ptr()), "process")
        return check(lib.pebble_decimator_process(self._h, _numSamples, _host_ptr(_in, _numSamples),
                                                  _host_ptr(_out, nout, True)), "process")

    def GetState(self):
        n = check(lib.pebble_decimator_state_size(self._h))
        buf = np.empty(max(n, 1), dtype=np.uint8)
        check(lib.pebble_decimator_get_state(self._h, buf.ctypes.data, n), "GetState")
        return buf[:n]

    def SetState(self, buf):
        check(lib.pebble_decimator_set_state(self._h, buf.ctypes.data, buf.size), "SetState")


class Mixer(_Handle):
    """Frequency shift by the recursive oscillator (pebblelib/mixer.cpp:5-81)."""
    _destroy = staticmethod(lib.pebble_mixer_destroy)

    def __init__(self, _sampleRate, _bufferSize):
        super().__init__(lib.pebble_mixer_create(int(_sampleRate), int(_bufferSize)), "Mixer")
        self._n = int(_bufferSize)
        self._out = None

    def setFrequency(self, f):
        lib.pebble_mixer_set_frequency(self._h, f)

    def processBlock(self, _in, n=None):
        """Returns `_in` itself when the frequency is 0 (mixer.cpp:51-53), else the step's own output buffer."""
        n = self._n if n is None else n
        if lib.pebble_mixer_is_passthrough(self._h) == 1:
            return _in
        if _is_cuda_tensor(_in):
            if self._out is None or not _is_cuda_tensor(self._out) or self._out.numel() < n:
                self._out = torch.empty(n, dtype=torch.complex128, device=_in.device)
            check(lib.pebble_mixer_process_device(self._h, n, _dev_ptr(_in, n), _dev_ptr(self._out, n), _stream_ptr()),
                  "processBlock")
            return self._out[:n]
        if self._out is None or _is_cuda_tensor(self._out) or self._out.size < n:
            self._out = np.empty(n, dtype=np.complex128)
        check(lib.pebble_mixer_process(self._h, n, _host_ptr(_in, n), _host_ptr(self._out, n, True)), "processBlock")
        return self._out[:n]


def _np_ptr(a, count, dtype, writable=False):
    if not isinstance(a, np.ndarray) or a.dtype != dtype or not a.flags.c_contiguous or a.size < count:
        raise PebbleError("expected a C-contiguous numpy %s array with >= %d elements" % (np.dtype(dtype), count))
    if writable and not a.flags.writeable:
        raise PebbleError("output array is read-only")
    return a.ctypes.data


class WindowFunction:
    """pebblelib/windowfunction.h: WINDOWTYPE numbering and table builder."""
    RECTANGULAR, HANNING, WELCH, PARZEN, BARTLETT, HAMMING, BLACKMAN2, BLACKMAN3, BLACKMAN4, EXPONENTIAL, RIEMANN, \
        BLACKMANHARRIS, BLACKMAN, NONE = range(1, 15)

    def __init__(self, _type, _numSamples):
        self.window = np.zeros(_numSamples)
        self.coherentGain = lib.pebble_window(_type, _numSamples, self.window.ctypes.data)


class FFT(_Handle):
    """pebblelib/fft.h:29-62 — fftParams / fftForward / fftInverse / fftSpectrum."""
    _destroy = staticmethod(lib.pebble_fft_destroy)

    def __init__(self):
        super().__init__(lib.pebble_fft_create(), "FFT")

    def fftParams(self, _fftSize, _dBCompensation, _sampleRate, _samplesPerBuffer, _windowType):
        check(lib.pebble_fft_params(self._h, _fftSize, _dBCompensation, _sampleRate, _samplesPerBuffer, _windowType),
              "fftParams")

    def getFFTSize(self):
        return check(lib.pebble_fft_size(self._h))

    def setBackend(self, backend):
        """0 = FFTfftw contract (default), 1 = CFft / Ooura contract (pebblelib/fftcute.cpp:183-277)."""
        check(lib.pebble_fft_set_backend(self._h, backend), "setBackend")

    def fftForward(self, inb, outb, numSamples):
        check(lib.pebble_fft_forward(self._h, _host_ptr(inb, numSamples), _host_ptr(outb, self.getFFTSize(), True),
                                     numSamples), "fftForward")

    def fftInverse(self, inb, outb, numSamples):
        check(lib.pebble_fft_inverse(self._h, _host_ptr(inb, numSamples), _host_ptr(outb, self.getFFTSize(), True),
                                     numSamples), "fftInverse")

    def fftSpectrum(self, inb, out, numSamples):
        """Returns the overload flag like the reference (fftw.cpp:70-81)."""
        return bool(check(lib.pebble_fft_spectrum(self._h, _host_ptr(inb, numSamples),
                                                  _np_ptr(out, self.getFFTSize(), np.float64, True), numSamples),
                          "fftSpectrum"))

    def fftSpectrumBatch(self, frames, numSamples, d_in, d_out_db, d_flags=None):
        """Device-resident batch: `frames` consecutive frames (torch CUDA tensors)."""
        return check(lib.pebble_fft_spectrum_batch_device(
            self._h, frames, numSamples, _dev_ptr(d_in, frames * numSamples), d_out_db.data_ptr(),
            d_flags.data_ptr() if d_flags is not None else None, _stream_ptr()), "fftSpectrumBatch")

    def GetState(self):
        buf = np.zeros(2 * self.getFFTSize())
        check(lib.pebble_fft_get_state(self._h, buf.ctypes.data, buf.size))
        return buf

    def SetState(self, buf):
        check(lib.pebble_fft_set_state(self._h, buf.ctypes.data, buf.size))


class CIir(_Handle):
    """pebblelib/iir.h: biquad (InitLP / InitHP / InitBP / InitBR + ProcessFilter on CPX)."""
    _destroy = staticmethod(lib.pebble_iir_destroy)
    LP, HP, BP, BR = range(4)

    def __init__(self, kind=None, f0=0.0, q=1.0, fs=1.0):
        super().__init__(lib.pebble_iir_create(), "CIir")
        if kind is not None:
            self.init(kind, f0, q, fs)

    def init(self, kind, f0, q, fs):
        check(lib.pebble_iir_init(self._h, kind, float(f0), float(q), float(fs)), "CIir.Init")

    def coefficients(self):
        c = np.zeros(5)
        check(lib.pebble_iir_get_coefficients(self._h, c.ctypes.data))
        return c

    def ProcessFilter(self, InLength, InBuf, OutBuf):
        if _is_cuda_tensor(InBuf):
            return check(lib.pebble_iir_process_device(self._h, InLength, _dev_ptr(InBuf, InLength), _dev_ptr(OutBuf, InLength),
                                                       _stream_ptr()), "ProcessFilter")
        return check(lib.pebble_iir_process(self._h, InLength, _host_ptr(InBuf, InLength), _host_ptr(OutBuf, InLength, True)),
                     "ProcessFilter")

    def GetState(self):
        w = np.zeros(4)
        check(lib.pebble_iir_get_state(self._h, w.ctypes.data))
        return w

    def SetState(self, w):
        w = np.ascontiguousarray(w, dtype=np.float64)
        check(lib.pebble_iir_set_state(self._h, w.ctypes.data))


class DCRemoval(CIir):
    """application/dcremoval.cpp: 10 Hz high-pass biquad, Q = 0.7071; a disabled step copies."""

    def __init__(self, sampleRate, bufferSize=0):
        super().__init__(CIir.HP, 10, 0.7071, sampleRate)
        self.enabled = False

    def enableStep(self, on):
        self.enabled = bool(on)

    def process(self, inb, n, out):
        if not self.enabled:
            out[:n] = inb[:n]
            return out
        self.ProcessFilter(n, inb, out)
        return out


def iqbalance(d_in, block_len, gainFactor=1.0, phaseFactor=0.0):
    """IQBalance::ProcessBlock (enabled) over consecutive blocks of block_len samples; CUDA tensor in, CUDA tensor out."""
    n = d_in.numel()
    out = torch.empty_like(d_in)
    check(lib.pebble_iqbalance_process_device(n, int(block_len), float(gainFactor), float(phaseFactor), d_in.data_ptr(), out.data_ptr(),
                                              _stream_ptr()), "iqbalance")
    return out


def fdEstimate(d_spectrum, spectrumSampleRate, bpLowFreq, bpHighFreq, mixerFreqs, squelchDb=-120.0):
    """SignalStrength::fdEstimate for one or many channels over one dB spectrum (CUDA float64 tensor).
    Returns [nch, 5]: peakDb, avgDb, snrDb, floorDb, open."""
    mix = np.ascontiguousarray(np.atleast_1d(mixerFreqs), dtype=np.float64)
    out = np.zeros((len(mix), 5))
    check(lib.pebble_sigstrength_fd_estimate_device(d_spectrum.data_ptr(), d_spectrum.numel(), int(spectrumSampleRate), float(bpLowFreq),
                                                    float(bpHighFreq), mix.ctypes.data, len(mix), float(squelchDb), out.ctypes.data,
                                                    _stream_ptr()), "fdEstimate")
    return out


class NoiseBlanker(_Handle):
    """application/noiseblanker.h: ProcessBlock (NB1) / ProcessBlock2 (NB2); both disabled at construction like the reference."""
    _destroy = staticmethod(lib.pebble_noiseblanker_destroy)

    def __init__(self, sampleRate, bufferSize):
        super().__init__(lib.pebble_noiseblanker_create(int(sampleRate), int(bufferSize)), "NoiseBlanker")
        self.numSamples = int(bufferSize)

    def setNbEnabled(self, b):
        check(lib.pebble_noiseblanker_set_nb_enabled(self._h, int(bool(b))), "setNbEnabled")

    def setNb2Enabled(self, b):
        check(lib.pebble_noiseblanker_set_nb2_enabled(self._h, int(bool(b))), "setNb2Enabled")

    def _run(self, which, inb):
        if lib.pebble_noiseblanker_enabled(self._h, which) != 1:
            return inb                                     # the reference returns `in` itself when the blanker is off
        n = len(inb)
        out = np.empty(n, dtype=np.complex128)
        check(lib.pebble_noiseblanker_process(self._h, which, n, _host_ptr(inb, n), out.ctypes.data), "NoiseBlanker")
        return out

    def ProcessBlock(self, inb):
        return self._run(1, inb)

    def ProcessBlock2(self, inb):
        return self._run(2, inb)


class RdsSync(_Handle):
    """The RDS block / group synchroniser alone (Demod_WFM::processNewRdsBit + checkBlock + getNextRdsGroupData,
    demod_wfm.cpp:584-788): host logic, fed with data bits."""
    _destroy = staticmethod(lib.pebble_rds_sync_destroy)

    def __init__(self):
        super().__init__(lib.pebble_rds_sync_create(), "RdsSync")

    def pushBits(self, bits):
        b = np.ascontiguousarray(bits, dtype=np.uint8)
        check(lib.pebble_rds_sync_push_bits(self._h, b.ctypes.data, len(b)), "pushBits")

    def getNextRdsGroupData(self):
        b = (ctypes.c_ushort * 4)()
        return tuple(b) if lib.pebble_rds_sync_next_group(self._h, b) == 1 else None


class NCO(_Handle):
    """pebblelib/nco.h: the oscillator primitive. setFrequency clamps to +-((sampleRate / 2) - 1) (nco.cpp:16-45)."""
    _destroy = staticmethod(lib.pebble_nco_destroy)

    def __init__(self, sampleRate, bufferSize):
        super().__init__(lib.pebble_nco_create(int(sampleRate), int(bufferSize)), "NCO")

    def setFrequency(self, f):
        return lib.pebble_nco_set_frequency(self._h, float(f))

    def frequency(self):
        return lib.pebble_nco_frequency(self._h)

    def genSingle(self, io, numSamples, dbGain, mix):
        if _is_cuda_tensor(io):
            return check(lib.pebble_nco_gen_single_device(self._h, _dev_ptr(io, numSamples), numSamples, dbGain, int(bool(mix)),
                                                          _stream_ptr()), "genSingle")
        return check(lib.pebble_nco_gen_single(self._h, _host_ptr(io, numSamples, True), numSamples, dbGain, int(bool(mix))), "genSingle")


def cpx_util(name, a, b=None, s=0.0):
    """CpxUtil helpers on CUDA tensors (complex128): copy, scale, mult, add, mag, sqrmag, clear, decimate."""
    n = a.numel()
    st = _stream_ptr()
    if name == "decimate":
        out = torch.zeros((n + int(s) - 1) // int(s), dtype=torch.complex128, device=a.device)
        check(lib.pebble_cpx_decimate_device(out.data_ptr(), a.data_ptr(), int(s), n, st), name)
        return out
    out = torch.empty_like(a)
    if name == "copy":
        check(lib.pebble_cpx_copy_device(out.data_ptr(), a.data_ptr(), n, st), name)
    elif name == "scale":
        check(lib.pebble_cpx_scale_device(out.data_ptr(), a.data_ptr(), float(s), n, st), name)
    elif name == "mult":
        check(lib.pebble_cpx_mult_device(out.data_ptr(), a.data_ptr(), b.data_ptr(), n, st), name)
    elif name == "add":
        check(lib.pebble_cpx_add_device(out.data_ptr(), a.data_ptr(), b.data_ptr(), n, st), name)
    elif name == "mag":
        check(lib.pebble_cpx_mag_device(out.data_ptr(), a.data_ptr(), n, st), name)
    elif name == "sqrmag":
        check(lib.pebble_cpx_sqrmag_device(out.data_ptr(), a.data_ptr(), n, st), name)
    elif name == "clear":
        check(lib.pebble_cpx_clear_device(out.data_ptr(), n, st), name)
    else:
        raise PebbleError("unknown CpxUtil op " + name)
    return out


class SignalSpectrum(_Handle):
    """application/signalspectrum.h: `unprocessed` + `zoomed` spectra behind the display-rate gate. now_ms = caller's clock."""
    _destroy = staticmethod(lib.pebble_spectrum_destroy)

    def __init__(self, sampleRate, hiResSampleRate, bufferSize, numSpectrumBins=4096, numHiResSpectrumBins=2048, updatesPerSecond=10):
        super().__init__(lib.pebble_spectrum_create(int(sampleRate), int(hiResSampleRate), int(bufferSize), int(numSpectrumBins),
                                                    int(numHiResSpectrumBins), int(updatesPerSecond)), "SignalSpectrum")

    def setHiRes(self, on):
        lib.pebble_spectrum_set_hires(self._h, int(bool(on)))

    def setUpdatesPerSec(self, u):
        lib.pebble_spectrum_set_updates_per_sec(self._h, int(u))

    def setSampleRate(self, fs, hires_fs):
        check(lib.pebble_spectrum_set_sample_rate(self._h, int(fs), int(hires_fs)), "setSampleRate")

    def unprocessed(self, inb, numSamples, now_ms):
        return check(lib.pebble_spectrum_unprocessed(self._h, _host_ptr(inb, numSamples), numSamples, int(now_ms)), "unprocessed")

    def zoomed(self, inb, numSamples, now_ms):
        return check(lib.pebble_spectrum_zoomed(self._h, _host_ptr(inb, numSamples), numSamples, int(now_ms)), "zoomed")

    def binCount(self):
        return check(lib.pebble_spectrum_bin_count(self._h))

    def getOverload(self):
        return bool(lib.pebble_spectrum_overload(self._h))

    def overrun(self):
        return lib.pebble_spectrum_overrun(self._h)

    def ackUpdate(self):
        lib.pebble_spectrum_ack_update(self._h)

    def getUnprocessed(self):
        n = self.binCount()
        p = lib.pebble_spectrum_get_unprocessed(self._h)
        return np.ctypeslib.as_array(ctypes.cast(p, ctypes.POINTER(ctypes.c_double)), shape=(n,)).copy()

    def getHiRes(self):
        n = check(lib.pebble_spectrum_hires_bin_count(self._h))
        p = lib.pebble_spectrum_get_hires(self._h)
        return np.ctypeslib.as_array(ctypes.cast(p, ctypes.POINTER(ctypes.c_double)), shape=(n,)).copy()


class CFir(_Handle):
    """pebblelib/fir.h:31-56."""
    _destroy = staticmethod(lib.pebble_cfir_destroy)

    def __init__(self):
        super().__init__(lib.pebble_cfir_create(), "CFir")

    def InitLPFilter(self, NumTaps, Scale, Astop, Fpass, Fstop, Fsamprate):
        return check(lib.pebble_cfir_init_lp(self._h, NumTaps, Scale, Astop, Fpass, Fstop, Fsamprate), "InitLPFilter")

    def InitHPFilter(self, NumTaps, Scale, Astop, Fpass, Fstop, Fsamprate):
        return check(lib.pebble_cfir_init_hp(self._h, NumTaps, Scale, Astop, Fpass, Fstop, Fsamprate), "InitHPFilter")

    def GetTaps(self):
        a = np.zeros(75)
        n = check(lib.pebble_cfir_get_taps(self._h, a.ctypes.data, 75))
        return a[:n].copy()

    def ProcessFilter(self, InLength, InBuf, OutBuf):
        if _is_cuda_tensor(InBuf):
            check(lib.pebble_cfir_process_device(self._h, InLength, _dev_ptr(InBuf, InLength), _dev_ptr(OutBuf, InLength),
                                                 _stream_ptr()), "ProcessFilter")
        else:
            check(lib.pebble_cfir_process(self._h, InLength, _host_ptr(InBuf, InLength), _host_ptr(OutBuf, InLength, True)),
                  "ProcessFilter")


class AGC(_Handle):
    """application/agc.h."""
    AGC_OFF, ACG_FAST, AGC_MED, AGC_SLOW, AGC_LONG = range(5)
    _destroy = staticmethod(lib.pebble_agc_destroy)

    def __init__(self, _sampleRate, _bufferSize):
        super().__init__(lib.pebble_agc_create(int(_sampleRate), int(_bufferSize)), "AGC")
        self.numSamples = int(_bufferSize)

    def setAgcMode(self, m, threshold):
        check(lib.pebble_agc_set_mode(self._h, m, threshold), "setAgcMode")

    def setAgcThreshold(self, g):
        check(lib.pebble_agc_set_threshold(self._h, g), "setAgcThreshold")

    def processBlock(self, pInData, n=None):
        """Returns the step's OWN output buffer (processstep.h:37-40): valid until the next call, not to be freed."""
        n = self.numSamples if n is None else n
        if getattr(self, "_out", None) is None or self._out.size < n:
            self._out = np.empty(n, dtype=np.complex128)
        check(lib.pebble_agc_process(self._h, n, _host_ptr(pInData, n), self._out.ctypes.data), "processBlock")
        return self._out[:n]


class Demod(_Handle):
    """application/demod.h — Demod::processBlock mode dispatch."""
    dmAM, dmSAM, dmFMN, dmFMM, dmFMS, dmDSB, dmLSB, dmUSB, dmCWL, dmCWU, dmDIGL, dmDIGU, dmNONE = range(13)
    _destroy = staticmethod(lib.pebble_demod_destroy)

    def __init__(self, _sampleRate, _bufferSize):
        super().__init__(lib.pebble_demod_create(int(_sampleRate), int(_bufferSize)), "Demod")

    def setDemodMode(self, _mode, _sourceSampleRate=0, _audioSampleRate=0):
        check(lib.pebble_demod_set_mode(self._h, _mode), "setDemodMode")

    def setNfmVariant(self, variant):
        check(lib.pebble_demod_set_nfm_variant(self._h, variant))

    def setWfmSampleRate(self, rate):
        """The second rate of Demod(sampleRate, wfmSampleRate, bufferSize) (demod.cpp:50-66)."""
        check(lib.pebble_demod_set_wfm_sample_rate(self._h, float(rate)), "setWfmSampleRate")

    def setBandwidth(self, bandwidth):
        check(lib.pebble_demod_set_bandwidth(self._h, bandwidth), "setBandwidth")

    def getStereoLock(self):
        """Demod_WFM::getStereoLock: the pilot lock decision of the last dmFMS block."""
        return bool(check(lib.pebble_demod_get_stereo_lock(self._h), "getStereoLock"))

    def setRds(self, on):
        """dmFMS: the RDS branch (demod_wfm.cpp:298-357) on (default, as in the reference) or off (audio path alone)."""
        check(lib.pebble_demod_set_rds(self._h, 1 if on else 0), "setRds")

    def getRdsRate(self):
        """m_RdsOutputRate: the sample rate of the RDS branch after its down-converter."""
        return float(lib.pebble_demod_get_rds_rate(self._h))

    def getNextRdsGroupData(self):
        """Demod_WFM::getNextRdsGroupData (demod_wfm.cpp:764-788): pops one queued group; the four 16-bit blocks when it
        differs from the last group returned, else None (queue empty or an unchanged group)."""
        b = (ctypes.c_ushort * 4)()
        return tuple(b) if lib.pebble_demod_get_next_rds_group(self._h, b) == 1 else None

    def getRdsTap(self, max_samples):
        """Test tap: the last dmFMS call's m_RdsRaw as the reference leaves it (:336-352)."""
        out = np.empty(max_samples, dtype=np.complex128)
        k = check(lib.pebble_demod_get_rds_tap(self._h, out.ctypes.data, max_samples), "getRdsTap")
        return out[:k].copy()

    def processBlock(self, inb, bufSize):
        """Like the reference, pass-through modes return the INPUT array itself (demod.cpp:127-138)."""
        if lib.pebble_demod_is_passthrough(self._h) == 1:
            return inb
        if getattr(self, "_out", None) is None or self._out.size < bufSize:
            self._out = np.empty(bufSize, dtype=np.complex128)   # the step owns `out` (processstep.h:37-40)
        check(lib.pebble_demod_process(self._h, bufSize, _host_ptr(inb, bufSize), self._out.ctypes.data), "processBlock")
        return self._out[:bufSize]


class CFractResampler(_Handle):
    """Fractional resampler (pebblelib/fractresampler.cpp:64-250)."""
    _destroy = staticmethod(lib.pebble_resampler_destroy)

    def __init__(self, MaxInputSize=2048, streams=1):
        super().__init__(lib.pebble_resampler_create(int(MaxInputSize)), "CFractResampler")   # = ctor + Init(MaxInputSize)
        self.streams = streams
        if streams != 1:
            check(lib.pebble_resampler_set_streams(self._h, streams), "set_streams")

    def OutCount(self, InLength, Rate, frame=0):
        return check(lib.pebble_resampler_out_count(self._h, InLength, frame, Rate), "OutCount")

    def Resample(self, InLength, Rate, pInBuf, pOutBuf=None, frame=0, clip=False):
        """Host numpy arrays: one reference call. CUDA tensors ([streams, n] rows or 1-D): `frame` > 0 = a batch of
        consecutive reference calls of `frame` samples. Returns (count, out)."""
        k = self.OutCount(InLength, Rate, frame)
        if _is_cuda_tensor(pInBuf):
            rows = pInBuf if pInBuf.dim() == 2 else pInBuf.view(1, -1)
            if pOutBuf is None:
                pOutBuf = torch.zeros(rows.shape[0], max(k, 1), dtype=torch.complex128, device=pInBuf.device)
            orows = pOutBuf if pOutBuf.dim() == 2 else pOutBuf.view(1, -1)
            r = check(lib.pebble_resampler_resample_device(self._h, InLength, frame, Rate, rows.data_ptr(), rows.stride(0),
                                                           orows.data_ptr(), orows.stride(0), int(bool(clip)), _stream_ptr()), "Resample")
            return r, pOutBuf
        if pOutBuf is None:
            pOutBuf = np.empty(max(k, 1), dtype=np.complex128)
        r = check(lib.pebble_resampler_resample(self._h, InLength, Rate, _host_ptr(pInBuf, InLength), _host_ptr(pOutBuf, k, True)),
                  "Resample")
        return r, pOutBuf

    def ResampleStereo16(self, InLength, Rate, d_in, gain, frame=0):
        k = self.OutCount(InLength, Rate, frame)
        rows = d_in if d_in.dim() == 2 else d_in.view(1, -1)
        out = torch.zeros(rows.shape[0], max(k, 1), 2, dtype=torch.int16, device=d_in.device)
        r = check(lib.pebble_resampler_resample_stereo16_device(self._h, InLength, frame, Rate, rows.data_ptr(), rows.stride(0),
                                                                out.data_ptr(), out.stride(0), float(gain), _stream_ptr()), "Resample")
        return r, out

    def GetState(self):
        n = check(lib.pebble_resampler_state_size(self._h))
        buf = np.empty(n, dtype=np.uint8)
        check(lib.pebble_resampler_get_state(self._h, buf.ctypes.data, n), "GetState")
        return buf

    def SetState(self, buf):
        check(lib.pebble_resampler_set_state(self._h, buf.ctypes.data, buf.size), "SetState")


def resampler_plan(float_time, rate, n):
    """Host-only planner: (output count, m_FloatTime afterwards, output times) of one Resample call."""
    te = ctypes.c_double(0.0)
    cnt = check(lib.pebble_resampler_plan(float(float_time), float(rate), int(n), ctypes.byref(te)), "plan")
    times = np.empty(max(cnt, 1))
    check(lib.pebble_resampler_plan_times(float(float_time), float(rate), int(n), times.ctypes.data, cnt), "plan_times")
    return cnt, te.value, times[:cnt]


class Receiver(_Handle):
    """The non-WFM branch of Receiver::processIQData (application/receiver.cpp:903-996) as one chain."""
    _destroy = staticmethod(lib.pebble_receiver_destroy)

    def __init__(self, sampleRate, framesPerBuffer=2048):
        super().__init__(lib.pebble_receiver_create(float(sampleRate), int(framesPerBuffer)), "Receiver")
        self.framesPerBuffer = framesPerBuffer
        self._fmt = IQ_CPX

    def setInputFormat(self, fmt, order=IQO_IQ, userIQGain=1.0, normalizeIQGain=1.0):
        """processIQData then takes raw device samples (DeviceInterfaceBase::normalizeIQ fused into the mixer load)."""
        check(lib.pebble_receiver_set_input_format(self._h, fmt, order, userIQGain, normalizeIQGain), "setInputFormat")
        self._fmt = fmt

    def setDataRate(self, maxBW):
        r = lib.pebble_receiver_set_data_rate(self._h, maxBW)
        if r <= 0:
            raise PebbleError("setDataRate: " + lib.pebble_last_error().decode())
        return r

    def setMixer(self, freq):
        lib.pebble_receiver_set_mixer(self._h, freq)

    def setFilter(self, lo, hi):
        return check(lib.pebble_receiver_set_filter(self._h, lo, hi), "setFilter")

    def setDemodMode(self, mode):
        check(lib.pebble_receiver_set_demod_mode(self._h, mode), "setDemodMode")

    def setDemodBandwidth(self, bw):
        check(lib.pebble_receiver_set_demod_bandwidth(self._h, bw), "setDemodBandwidth")

    def setAgc(self, mode, threshold):
        check(lib.pebble_receiver_set_agc(self._h, mode, threshold), "setAgc")

    def setGainStages(self, stages):
        check(lib.pebble_receiver_set_gain_stages(self._h, stages))

    def setAudio(self, audioRate, clip=True):
        """Resample the demodulated frames to audioRate (receiver.cpp:994-1002) and clip like the audio sink."""
        check(lib.pebble_receiver_set_audio(self._h, int(audioRate), int(bool(clip))), "setAudio")

    def outCount(self, n):
        return check(lib.pebble_receiver_out_count(self._h, n))

    def wfmRate(self):
        """m_demodWfmSampleRate (receiver.cpp:217): the rate the WFM branch demodulates at."""
        return float(lib.pebble_receiver_get_wfm_rate(self._h))

    def getStereoLock(self):
        """What Demod::fmStereo reads after every block (demod.cpp:201-205)."""
        return bool(check(lib.pebble_demod_get_stereo_lock(lib.pebble_receiver_demod(self._h)), "getStereoLock"))

    def getNextRdsGroupData(self):
        b = (ctypes.c_ushort * 4)()
        return tuple(b) if lib.pebble_demod_get_next_rds_group(lib.pebble_receiver_demod(self._h), b) == 1 else None

    def setProfiling(self, on):
        check(lib.pebble_receiver_set_profiling(self._h, int(bool(on))))

    def getProfile(self):
        """({"downconvert": ms, "fastfir": ms, "agc": ms, "demod": ms}, calls) summed since setProfiling(True)."""
        ms = (ctypes.c_double * 4)()
        calls = ctypes.c_int64(0)
        check(lib.pebble_receiver_get_profile(self._h, ms, ctypes.byref(calls)), "getProfile")
        return dict(zip(("downconvert", "fastfir", "agc", "demod"), list(ms))), calls.value

    def processIQData(self, inb, numSamples, out=None):
        """Host arrays (numpy) or device tensors (torch). Returns (count, out)."""
        k = self.outCount(numSamples)
        if _is_cuda_tensor(inb):
            if out is None:
                out = torch.empty(max(k, 1), dtype=torch.complex128, device=inb.device)
            r = check(lib.pebble_receiver_process_device(self._h, numSamples, _raw_ptr(inb, numSamples, self._fmt),
                                                         _dev_ptr(out, k), _stream_ptr()), "processIQData")
            return r, out
        if out is None:
            out = np.empty(max(k, 1), dtype=np.complex128)
        r = check(lib.pebble_receiver_process(self._h, numSamples, _raw_ptr(inb, numSamples, self._fmt), _host_ptr(out, k, True)),
                  "processIQData")
        return r, out

    def processPinned(self, pinned_in, numSamples, pinned_out):
        """pinned_in/out: torch CPU tensors in page-locked memory; enqueues H2D, chain, D2H on the current stream."""
        return check(lib.pebble_receiver_process_pinned(self._h, numSamples, pinned_in.data_ptr(), pinned_out.data_ptr(),
                                                        _stream_ptr()), "processPinned")


class Channelizer(_Handle):
    """nch receiver chains (receiver.cpp:903-996) over one wideband capture — BASELINE config 4. Device tensors only."""
    _destroy = staticmethod(lib.pebble_channelizer_destroy)

    def __init__(self, nchannels, sampleRate, maxBW):
        super().__init__(lib.pebble_channelizer_create(int(nchannels), float(sampleRate), float(maxBW)), "Channelizer")
        self.nchannels = int(nchannels)
        self.outRate = lib.pebble_channelizer_out_rate(self._h)
        self.numStages = check(lib.pebble_channelizer_num_stages(self._h))
        self._fmt = IQ_CPX

    def setInputFormat(self, fmt, order=IQO_IQ, userIQGain=1.0, normalizeIQGain=1.0):
        """The capture then holds raw device samples of this format (normalizeIQ fused into the front end's load)."""
        check(lib.pebble_channelizer_set_input_format(self._h, fmt, order, userIQGain, normalizeIQGain), "setInputFormat")
        self._fmt = fmt

    def frontPath(self):
        """1 = polyphase-DFT front end ran the last call (channels on the grid fs / 2^S), 0 = per-channel kernel, -1 = none yet."""
        return lib.pebble_channelizer_front_path(self._h)

    def setFrequency(self, ch, freq):
        check(lib.pebble_channelizer_set_frequency(self._h, ch, freq), "setFrequency")

    def setFilter(self, lo, hi):
        return check(lib.pebble_channelizer_set_filter(self._h, lo, hi), "setFilter")

    def setDemodMode(self, ch, mode):
        check(lib.pebble_channelizer_set_demod_mode(self._h, ch, mode), "setDemodMode")

    def setAmBandwidth(self, bw):
        check(lib.pebble_channelizer_set_am_bandwidth(self._h, bw), "setAmBandwidth")

    def setAgcThreshold(self, thr):
        check(lib.pebble_channelizer_set_agc_threshold(self._h, thr))

    def outCount(self, n):
        return check(lib.pebble_channelizer_out_count(self._h, n))

    def setSquelch(self, open_mask=None):
        """Per-channel squelch decision (receiver.cpp:955-965): 0 = the channel yields zeros and its demodulator state does not
        advance. None: every channel open."""
        if open_mask is None:
            return check(lib.pebble_channelizer_set_squelch(self._h, None, 0), "setSquelch")
        m = np.ascontiguousarray(open_mask, dtype=np.uint8)
        return check(lib.pebble_channelizer_set_squelch(self._h, m.ctypes.data, int(m.size)), "setSquelch")

    def squelchFromSpectrum(self, d_spectrum, spectrumSampleRate, squelchDb):
        """fdEstimate for every channel over one wideband dB spectrum (CUDA float64 tensor); returns the number of open channels."""
        return check(lib.pebble_channelizer_squelch_from_spectrum(self._h, d_spectrum.data_ptr(), d_spectrum.numel(), int(spectrumSampleRate),
                                                                  float(squelchDb), _stream_ptr()), "squelchFromSpectrum")

    def downConvert(self, d_in, n, d_out):
        """d_out: [nchannels, >= n >> S] complex128 CUDA tensor (rows contiguous)."""
        assert d_out.dim() == 2 and d_out.shape[0] == self.nchannels and d_out.stride(1) == 1
        return check(lib.pebble_channelizer_downconvert_device(self._h, n, _raw_ptr(d_in, n, self._fmt), d_out.data_ptr(),
                                                               d_out.stride(0), _stream_ptr()), "downConvert")

    def process(self, d_in, n, d_out):
        """d_out: [nchannels, >= n >> S] complex128 CUDA tensor. Returns demodulated samples per channel."""
        assert d_out.dim() == 2 and d_out.shape[0] == self.nchannels and d_out.stride(1) == 1
        return check(lib.pebble_channelizer_process_device(self._h, n, _raw_ptr(d_in, n, self._fmt), d_out.data_ptr(),
                                                           d_out.stride(0), _stream_ptr()), "process")
