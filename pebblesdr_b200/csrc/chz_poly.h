// chz_poly.h — polyphase-DFT front end of the batched channelizer (chz_poly.cu): internal interface used by the batched
// form of CDownConvert (downconvert.cu: pebble_downconvert_process_channels_device).
#pragma once
#include "lib_common.h"
#include "dc_types.h"
#include <vector>

struct ChzPoly;
ChzPoly *chz_poly_create();
void chz_poly_destroy(ChzPoly *p);
// forget the plan and the carried input history (SetDataRate / set_channels)
void chz_poly_invalidate(ChzPoly *p);
// 1: the channels sit on the uniform grid fs / 2^S and the polyphase path applies; 0: it does not; < 0: error
int chz_poly_prepare(ChzPoly *p, const DcChain &chain, const std::vector<DcChan> &chans);
bool chz_poly_active(const ChzPoly *p);
// n input samples (multiple of 2^S) of format `fmt` -> rows [channel][n >> S] of d_out, out_stride cpx apart. chans[k].phase_* =
// channel k's oscillator phase before sample 0 of this call. Returns n >> S.
int chz_poly_process(ChzPoly *p, const DcChain &chain, const DcChain *d_chain, const std::vector<DcChan> &chans, long long stream_pos,
                     int fmt, double in_scale, long long n, const void *d_in, cpx *d_out, long long out_stride, cudaStream_t st);
