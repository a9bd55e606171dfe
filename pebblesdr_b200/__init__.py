"""pebblesdr_b200 — PebbleSDR's receiver IQ hot path (DownConvert -> FastFIR -> demod, SignalSpectrum)
as hand-written sm_100a CUDA kernels behind a C ABI (include/pebble_cuda.h). The Python classes in
`ops` mirror the reference's operator surface (same names, arguments and return conventions) so the
parity tests read like calls into pebblelib; all arithmetic happens in libpebblecuda.so on the GPU."""
from ._lib import lib, PebbleError, LIB_PATH  # noqa: F401
from .ops import CFastFIR, CDownConvert, FFT, WindowFunction, CFir, AGC, Demod, Receiver, Channelizer, Decimator, Mixer, decimator_plan  # noqa: F401
from .ops import CFractResampler, resampler_plan, SignalSpectrum, NCO, cpx_util, CIir, DCRemoval, iqbalance, fdEstimate, NoiseBlanker, RdsSync  # noqa: F401
from .ops import normalizeIQ, IQ_CPX, IQ_CS16, IQ_CU8, IQ_CS8, IQ_CF32, IQO_IQ, IQO_QI, IQO_IONLY, IQO_QONLY  # noqa: F401
