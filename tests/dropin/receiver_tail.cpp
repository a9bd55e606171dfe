// tests/dropin/receiver_tail.cpp — Receiver::processIQData's non-WFM branch (application/receiver.cpp:826, 903-1007) written
// against include/pebble_dropin.hpp: the same statements, in the same order, with the reference's call shape — the device
// plugin's consumer thread hands over framesPerBuffer = 2048 CPX samples per call (quint16 count, device_interfaces.h:32).
// Built by g++ against libpebblecuda.so and run by tests/test_gpu_dropin.py, which checks the audio it writes against the
// oracle chain. It also times every call (SURVEY 7.3 "drop-in latency"): once through the per-operator classes (each
// operator call is its own H2D + kernel + D2H) and once through the fused pebble_receiver_* entry.
//
// usage: receiver_tail <in.c128> <out.c128> <mode: am|usb|fmn> [calls_for_latency]
#include "pebble_dropin.hpp"
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <vector>

namespace DB { inline double dBToAmplitude(double db) { return pow(10, db / 20.0); } }   // pebblelib/db.h:60-62

class Receiver {
public:
    Receiver(quint32 sampleRate, quint32 framesPerBuffer, const char *mode, std::vector<CPX> *audio)
        : m_sampleRate(sampleRate), m_framesPerBuffer(framesPerBuffer), m_audioOutRate(48000), m_sampleBufLen(0), m_audio(audio) {
        // Receiver::turnPowerOn, receiver.cpp:116-281 (only what the branch below touches)
        m_demodSampleRate = (quint32)m_downConvert1.SetDataRate(m_sampleRate, 15000.0);            // :196
        m_downConvert1.SetFrequency(1.2e6);                                                         // mixerChanged, :640-655
        m_signalSpectrum = new SignalSpectrum(m_sampleRate, m_demodSampleRate, m_framesPerBuffer, 4096, 2048, 10);
        m_signalSpectrum->setHiRes(true);
        m_bpFilter = new BandPassFilter(m_demodSampleRate, m_framesPerBuffer);
        m_agc = new AGC(m_demodSampleRate, m_framesPerBuffer);
        m_agc->setAgcMode(AGC::AGC_OFF, 20);
        m_demod = new Demod(m_demodSampleRate, m_framesPerBuffer);
        const bool am = mode[0] == 'a', fmn = mode[0] == 'f';
        m_demod->setDemodMode(am ? DeviceInterface::dmAM : fmn ? DeviceInterface::dmFMN : DeviceInterface::dmUSB, m_demodSampleRate, m_audioOutRate);
        if (am) { m_bpFilter->setBandPass(-5000, 5000); m_demod->setBandwidth(10000.0); }           // filterChanged, :658-664
        else if (fmn) m_bpFilter->setBandPass(-7500, 7500);
        else m_bpFilter->setBandPass(300, 3000);
        m_fractResampler.Init(m_framesPerBuffer);
        m_workingBuf = pebble_dropin::memalign(m_framesPerBuffer * 2);
        m_sampleBuf = pebble_dropin::memalign(m_framesPerBuffer * 2);
        m_audioBuf = pebble_dropin::memalign(m_framesPerBuffer * 2);
    }
    ~Receiver() {
        delete m_signalSpectrum; delete m_bpFilter; delete m_agc; delete m_demod;
        free(m_workingBuf); free(m_sampleBuf); free(m_audioBuf);
    }
    // receiver.cpp:758-1009, non-WFM branch
    void processIQData(CPX *in, quint16 numSamples) {
        CPX *nextStep = in;
        int numStepSamples = numSamples;
        double resampRate;
        m_signalSpectrum->unprocessed(nextStep, numSamples);                                        // :826
        numStepSamples = m_downConvert1.ProcessData(numStepSamples, nextStep, m_workingBuf);        // :908
        for (int i = 0; i < numStepSamples; i++) m_sampleBuf[m_sampleBufLen++] = m_workingBuf[i];  // :922-924
        if (m_sampleBufLen < (int)m_framesPerBuffer) return;                                        // :927-928
        numStepSamples = m_framesPerBuffer;
        m_sampleBufLen = 0;
        nextStep = m_sampleBuf;
        quint32 decimationLoss = m_downConvert1.NumStages();                                        // :934 (decBy2Stages of the active decimator)
        scaleCPX(nextStep, nextStep, DB::dBToAmplitude(decimationLoss * 2), numStepSamples);        // :937
        m_signalSpectrum->zoomed(nextStep, numStepSamples);                                         // :942
        nextStep = m_bpFilter->process(nextStep, numStepSamples);                                   // :950
        nextStep = m_agc->processBlock(nextStep);                                                   // :983
        nextStep = m_demod->processBlock(nextStep, numStepSamples);                                 // :987
        resampRate = (m_demodSampleRate * 1.0) / (m_audioOutRate * 1.0);                            // :994
        if (resampRate != 1)
            numStepSamples = m_fractResampler.Resample(numStepSamples, resampRate, nextStep, m_audioBuf);   // :1000-1001
        else
            copyCPX(m_audioBuf, nextStep, numStepSamples);
        processAudioData(m_audioBuf, numStepSamples);                                               // :1007
    }
    void processAudioData(CPX *in, quint16 numSamples) { m_audio->insert(m_audio->end(), in, in + numSamples); }   // :1030-1034 -> sink
private:
    quint32 m_sampleRate, m_framesPerBuffer, m_demodSampleRate, m_audioOutRate;
    CDownConvert m_downConvert1;
    SignalSpectrum *m_signalSpectrum;
    BandPassFilter *m_bpFilter;
    AGC *m_agc;
    Demod *m_demod;
    CFractResampler m_fractResampler;
    CPX *m_workingBuf, *m_sampleBuf, *m_audioBuf;
    int m_sampleBufLen;
    std::vector<CPX> *m_audio;
};

static void stats(std::vector<double> &us, const char *name) {
    std::sort(us.begin(), us.end());
    printf("\"%s\": {\"calls\": %zu, \"median_us\": %.2f, \"p99_us\": %.2f, \"max_us\": %.2f}", name, us.size(), us[us.size() / 2],
           us[(size_t)(us.size() * 0.99)], us.back());
}

int main(int argc, char **argv) {
    if (argc < 4) { fprintf(stderr, "usage: receiver_tail in.c128 out.c128 am|usb|fmn [latency_calls]\n"); return 2; }
    FILE *f = fopen(argv[1], "rb");
    if (!f) { perror(argv[1]); return 2; }
    fseek(f, 0, SEEK_END);
    const size_t n = (size_t)ftell(f) / sizeof(CPX);
    fseek(f, 0, SEEK_SET);
    std::vector<CPX> x(n);
    if (fread(x.data(), sizeof(CPX), n, f) != n) return 2;
    fclose(f);
    const quint16 fpb = 2048;
    std::vector<CPX> audio;
    std::vector<double> us;
    {
        Receiver rx(10000000, fpb, argv[3], &audio);
        for (size_t i = 0; i + fpb <= n; i += fpb) {
            const auto t0 = std::chrono::steady_clock::now();
            rx.processIQData(x.data() + i, fpb);
            us.push_back(std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count());
        }
    }
    f = fopen(argv[2], "wb");
    fwrite(audio.data(), sizeof(CPX), audio.size(), f);
    fclose(f);
    // the fused entry with the same call shape (one H2D + chain + D2H per call)
    const size_t lat_calls = argc > 4 ? (size_t)atol(argv[4]) : 2000;
    std::vector<double> us2;
    pebble_receiver_t *r = pebble_receiver_create(10e6, fpb);
    pebble_receiver_set_data_rate(r, 15000.0);
    pebble_receiver_set_mixer(r, 1.2e6);
    pebble_receiver_set_filter(r, -5000.0, 5000.0);
    pebble_receiver_set_demod_mode(r, 0);
    pebble_receiver_set_agc(r, 0, 20);
    pebble_receiver_set_demod_bandwidth(r, 10000.0);
    pebble_receiver_set_audio(r, 48000, 0);
    std::vector<CPX> out(4 * fpb);
    for (size_t c = 0; c < lat_calls; ++c) {
        const size_t i = (c * fpb) % (n - fpb);
        const auto t0 = std::chrono::steady_clock::now();
        pebble_receiver_process(r, fpb, reinterpret_cast<const double *>(x.data() + i), reinterpret_cast<double *>(out.data()));
        us2.push_back(std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count());
    }
    pebble_receiver_destroy(r);
    printf("{\"audio_samples\": %zu, ", audio.size());
    stats(us, "per_operator_classes");
    printf(", ");
    stats(us2, "fused_pebble_receiver_process");
    printf("}\n");
    return 0;
}
