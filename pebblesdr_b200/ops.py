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
            return check(lib.pebble_downconvert_process_device(self._h, InLength, _dev_ptr(pInData, InLength),
                                                               _dev_ptr(pOutData, nout), _stream_ptr()), "ProcessData")
        return check(lib.pebble_downconvert_process(self._h, InLength, _host_ptr(pInData, InLength),
                                                    _host_ptr(pOutData, nout, True)), "ProcessData")

    def GetState(self):
        n = check(lib.pebble_downconvert_state_size(self._h))
        buf = np.empty(n, dtype=np.uint8)
        check(lib.pebble_downconvert_get_state(self._h, buf.ctypes.data, n), "GetState")
        return buf

    def SetState(self, buf):
        check(lib.pebble_downconvert_set_state(self._h, buf.ctypes.data, buf.size), "SetState")
