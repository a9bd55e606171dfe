"""oracle/cpu_chain.py — TEST INFRASTRUCTURE ONLY (checker and timed CPU baseline; never the product path).

The reference's CPU receiver chain for BASELINE config 2, ordered as Receiver::processIQData's non-WFM branch
(application/receiver.cpp:903-996):

    CDownConvert::ProcessData (:908) -> m_sampleBuf re-framing to framesPerBuffer (:922-931) -> gain restore
    10^(2*stages/20) (:935-938) -> BandPassFilter/CFastFIR::ProcessData (:950) -> AGC_OFF gain (agc.cpp:86-94)
    -> Demod (USB: return in, demod.cpp:127-138; AM: Demod_AM::processBlockFiltered, demod_am.cpp:40-64)

kind "reference": every step is the REAL reference: CDownConvert and CFastFIR from oracle/_ref/libpebbleref.so
(unmodified pebblelib + FFTW 3.3.4), AGC, Demod_AM and DeviceInterfaceBase::normalizeIQ from
oracle/_ref/libpebbleref_app.so (unmodified application/agc.cpp, demod/demod_am.cpp and the normalizeIQ bodies of
pebblelib/deviceinterfacebase.cpp compiled against empty Qt stand-ins, oracle/build_ref.sh step 5).
kind "port": everything from oracle/pebble_oracle.c (used when oracle/_ref is absent).

Raw device samples: run_raw() starts from interleaved qint16 I,Q (CPX16, pebblelib/cpx.h:65-74) as a device plugin
receives them and converts each call's block with normalizeIQ (deviceinterfacebase.cpp:726-762) before
CDownConvert — the host-side work the CUDA path fuses into its mixer load.
"""
import os
import threading
import time

import numpy as np

from . import port, ref, ref_app

CALL = 32768  # samples per CDownConvert call (multiple of 128; generic HB buffers cap it, downconvert.cpp:60,345)


def kind():
    return "reference" if (ref.available() and ref_app.available()) else "port"


class CpuChain:
    def __init__(self, rate=10e6, fpb=2048, max_bw=15000.0, mixer=1.2e6, lo=-5000.0, hi=5000.0, demod_mode=0,
                 agc_threshold=20, am_bandwidth=10000.0, use_ref=None):
        self.use_ref = (ref.available() and ref_app.available()) if use_ref is None else use_ref
        self.fpb = fpb
        self.demod_mode = demod_mode
        if self.use_ref:
            self.dc = ref.CDownConvert()
            self.ff = ref.CFastFIR()
            self._dc_process = ref.lib().ref_downconvert_process
            self._ff_process = ref.lib().ref_fastfir_process
        else:
            self.dc = port.CDownConvert()
            self.ff = port.CFastFIR()
            self._dc_process = port.lib().orc_downconvert_process
            self._ff_process = port.lib().orc_fastfir_process
        self.demod_rate = self.dc.SetDataRate(rate, max_bw)
        self.dc.SetFrequency(mixer)
        self.ff.SetupParameters(lo, hi, 0.0, self.demod_rate)
        self.stages = int(round(np.log2(rate / self.demod_rate)))
        self.gain = 10.0 ** ((self.stages * 2) / 20.0)             # receiver.cpp:935-938
        if self.use_ref:
            self.agc = ref_app.AGC(self.demod_rate, fpb)              # AGC_OFF: out = m_manualIFGainAmp * in (agc.cpp:86-94)
            self.agc.setAgcMode(0, agc_threshold)
            self._agc = self.agc.processBlock
            self.demod = ref_app.Demod_AM(self.demod_rate, fpb)
            if demod_mode == 0:
                self.demod.setBandwidth(am_bandwidth)
            self._am = self.demod.processBlockFiltered
            self._norm = ref_app.lib().ref_normalize_iq
        else:
            self.agc = port.AGC(self.demod_rate)
            self.agc.setAgcMode(0, agc_threshold)
            self._agc = lambda y: self.agc.processBlock(np.ascontiguousarray(y), len(y))
            self.demod = port.Demod(self.demod_rate)
            if demod_mode == 0:
                self.demod.setBandwidth(am_bandwidth)
            self._am = lambda y: self.demod.am(np.ascontiguousarray(y), len(y))
            self._norm = port.lib().orc_normalize_iq
        self._conv = np.zeros(CALL, dtype=np.complex128)
        self.sample_buf = np.zeros(CALL + 2 * fpb, dtype=np.complex128)   # m_sampleBuf
        self.fill = 0
        self._dec = np.zeros(CALL, dtype=np.complex128)
        self._bp = np.zeros(fpb, dtype=np.complex128)

    def process(self, scratch, n, out):
        """scratch: n input samples (CLOBBERED, like the reference, downconvert.cpp:309-310). n a multiple of 128
        and <= CALL. Appends demodulated samples to out (list) and returns their count."""
        k = self._dc_process(self.dc.h, n, scratch.ctypes.data, self._dec.ctypes.data)
        self.sample_buf[self.fill:self.fill + k] = self._dec[:k]
        self.fill += k
        produced = 0
        while self.fill >= self.fpb:
            frame = self.sample_buf[:self.fpb] * self.gain
            self._ff_process(self.ff.h, self.fpb, frame.ctypes.data, self._bp.ctypes.data)
            y = self._agc(self._bp)
            if self.demod_mode == 0:
                y = self._am(y)
            out.append(y.copy())
            produced += self.fpb
            rem = self.fill - self.fpb
            self.sample_buf[:rem] = self.sample_buf[self.fpb:self.fill]
            self.fill = rem
        return produced

    def run(self, buf):
        """Whole buffer (clobbered) in CALL-sized calls; returns the concatenated demodulated stream."""
        out = []
        for i in range(0, len(buf), CALL):
            n = min(CALL, len(buf) - i)
            self.process(buf[i:i + n], n, out)
        return np.concatenate(out) if out else np.zeros(0, dtype=np.complex128)


    def run_raw(self, raw16):
        """raw16: interleaved qint16 I,Q (2 per sample). normalizeIQ (unit gains, IQ order) then the chain, per call."""
        out = []
        n_total = raw16.size // 2
        for i in range(0, n_total, CALL):
            n = min(CALL, n_total - i)
            self._norm(2, 0, 1.0, 1.0, 0, n, raw16[2 * i:].ctypes.data, self._conv.ctypes.data)   # fmt 2 = CPX16
            self.process(self._conv, n, out)
        return np.concatenate(out) if out else np.zeros(0, dtype=np.complex128)


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


BUF_CAP = 1 << 22   # samples per thread-private input buffer (64 MiB); longer steps re-fill it (untimed) and go again


def timed_throughput(make_input, samples_per_thread, threads, steps=1, warmup=0, chain_kwargs=None, raw16=False):
    """One independent receiver (own CpuChain, own input buffer) per thread — how N reference receivers would run on
    N cores; the reference itself is single-threaded per receiver. A step pushes `samples_per_thread` samples
    through every receiver as passes over a BUF_CAP-sample buffer; because CDownConvert clobbers its input, the
    buffer is re-filled before each pass and that copy is NOT timed (the device plugin hands the chain a fresh
    buffer). Step time = max over threads of the time spent inside the chain, all threads running concurrently.
    raw16=True: make_input returns interleaved qint16 I,Q and every pass runs normalizeIQ + chain (run_raw); nothing is
    clobbered, so there is no re-fill.
    Returns (list of per-step seconds, samples per step over all threads). ctypes releases the GIL inside the C calls."""
    chain_kwargs = chain_kwargs or {}
    per_pass = min(BUF_CAP, samples_per_thread) // CALL * CALL
    passes = max(1, samples_per_thread // per_pass)
    src = make_input(per_pass)
    chains = [CpuChain(**chain_kwargs) for _ in range(threads)]
    bufs = [src.copy() if raw16 else np.empty_like(src) for _ in range(threads)]
    times = []
    for step in range(warmup + steps):
        busy = [0.0] * threads
        start = threading.Barrier(threads)

        def work(i):
            start.wait()
            for _ in range(passes):
                if raw16:
                    t0 = time.perf_counter()
                    chains[i].run_raw(bufs[i])
                else:
                    np.copyto(bufs[i], src)
                    t0 = time.perf_counter()
                    chains[i].run(bufs[i])
                busy[i] += time.perf_counter() - t0

        ts = [threading.Thread(target=work, args=(i,)) for i in range(threads)]
        for t in ts:
            t.start()
        for t in ts:
            t.join()
        if step >= warmup:
            times.append(max(busy))
    return times, per_pass * passes * threads


class ChannelChain:
    """One channel of BASELINE config 4 on the CPU, every step the COMPILED reference: CDownConvert built with
    MAX_DECSTAGES 12 (oracle/build_ref.sh step 6; the stock build overflows on this 10-stage chain) ->
    gain restore 10^(2*stages/20) (receiver.cpp:935-938) -> CFastFIR(lo, hi) -> AGC_OFF(threshold) -> Demod_NFM::processBlockNCO
    (mode 2) / Demod_AM::processBlockFiltered (mode 0) / pass-through, once per 2048-sample frame (receiver.cpp:903-996).
    Fed in 65536-sample calls: the 35-tap last stage needs >= 2*34 samples per call (downconvert.cpp:387-390)."""
    CALL = 65536

    def __init__(self, rate, max_bw, freq, mode, lo=-7500.0, hi=7500.0, agc_threshold=20, fpb=2048):
        self.dc = ref_app.CDownConvert12()
        self.demod_rate = self.dc.SetDataRate(rate, max_bw)
        self.dc.SetFrequency(freq)
        self.S = int(round(np.log2(rate / self.demod_rate)))
        self.gain = 10.0 ** ((self.S * 2) / 20.0)
        self.ff = ref.CFastFIR()
        self.ff.SetupParameters(lo, hi, 0.0, self.demod_rate)
        self.agc = ref_app.AGC(self.demod_rate, fpb)
        self.agc.setAgcMode(0, agc_threshold)
        self.mode, self.fpb = mode, fpb
        self.dem = ref_app.Demod_NFM(self.demod_rate, fpb) if mode == 2 else ref_app.Demod_AM(self.demod_rate, fpb) if mode == 0 else None
        self.pend = np.zeros(0, dtype=np.complex128)
        self._bp = np.zeros(fpb, dtype=np.complex128)

    def run(self, x):
        out = []
        for i in range(0, len(x), self.CALL):
            n = min(self.CALL, len(x) - i)
            dec = np.zeros(n >> self.S, dtype=np.complex128)
            assert self.dc.ProcessData(n, x[i:i + n], dec) == n >> self.S
            self.pend = np.concatenate([self.pend, dec * self.gain])
            while len(self.pend) >= self.fpb:
                self.ff.ProcessData(self.fpb, np.ascontiguousarray(self.pend[:self.fpb]), self._bp)
                y = self.agc.processBlock(self._bp)
                if self.mode == 2:
                    y = self.dem.processBlockNCO(y)
                elif self.mode == 0:
                    y = self.dem.processBlockFiltered(y)
                out.append(np.array(y, copy=True))
                self.pend = self.pend[self.fpb:]
        return np.concatenate(out) if out else np.zeros(0, dtype=np.complex128)


def channel_chain_available():
    return ref.available() and ref_app.available() and ref_app.dec12_available()
