// demod_internal.h — handle layouts shared between demod.cu and receiver.cu (library-internal).
#pragma once
#include "lib_common.h"
#include "common.cuh"
#include "host_design.h"

struct pebble_cfir {
    int ntaps = 0;
    double fs = 0.0;
    double hI[75], hQ[75];
    double *d_hI = nullptr, *d_hQ = nullptr;
    cpx *d_hist[2] = {nullptr, nullptr};
    int cur = 0;
    cudaStream_t stream = nullptr;
    pebble::Staging stg;
};
int cfir_alloc(pebble_cfir *f);
void cfir_release(pebble_cfir *f);
int cfir_upload(pebble_cfir *f);   // after hI/hQ/ntaps changed: uploads taps and zeroes the history
// dup_input: the caller guarantees re == im for every input sample (lets equal I/Q tap sets take the real-only kernel)
int cfir_run_device(pebble_cfir *f, int n, const cpx *d_in, cpx *d_out, cudaStream_t st, bool dup_input = false);

struct pebble_agc;
struct pebble_demod;
// device-pointer forms used by the fused receiver chain
// frame: samples per reference Demod::processBlock call when n covers several (the receiver chain: framesPerBuffer); 0 = n
int demod_run_device(pebble_demod *h, int n, const cpx *d_in, cpx *d_out, cudaStream_t st, int *passthrough, int frame);
int agc_run_device(pebble_agc *h, int n, const cpx *d_in, cpx *d_out, cudaStream_t st);
void *demod_am_expect_tiles(pebble_demod *h, int n, double *alpha);
// fastfir.cu: the next process call also writes, per 1024-sample output block b, tile_map[b] = {alpha^1024,
// sum_k alpha^(1023-k) |y_k|} (two doubles), the AM detector's affine map of that block. One call only.
struct pebble_fastfir;
int fastfir_am_aggregate_next(pebble_fastfir *h, void *tile_map, double alpha, int n);
void demod_am_cancel_tiles(pebble_demod *h);
int demod_mode(const pebble_demod *h);
double agc_manual_gain(const pebble_agc *h);
int agc_is_off(const pebble_agc *h);

// rds.cu: the RDS branch of Demod_WFM::processDataStereo (demod_wfm.cpp:298-357) and its block / group synchroniser
struct RdsBranch;
RdsBranch *rds_create();
void rds_destroy(RdsBranch *r);
int rds_setup(RdsBranch *r, double wfm_rate);                                       // setSampleRate's RDS part + initRds
double rds_rate(const RdsBranch *r);
int rds_process(RdsBranch *r, int n, const cpx *d_cpx_raw_fm, cudaStream_t st);     // returns the RDS-rate sample count
int rds_next_group(RdsBranch *r, unsigned short *blocks4);                          // getNextRdsGroupData
int rds_tap(RdsBranch *r, double *out, int max_samples);
