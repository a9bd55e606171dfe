"""ctypes binding of libpebblecuda.so (include/pebble_cuda.h). The library is the product; this file only
declares signatures. There is no Python/CPU fallback: if the shared library is missing, import fails."""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libpebblecuda.so")

c_int, c_dbl, c_vp, c_i64, c_u64 = ctypes.c_int, ctypes.c_double, ctypes.c_void_p, ctypes.c_int64, ctypes.c_uint64

# name -> (restype, argtypes); every symbol include/pebble_cuda.h declares
SIGNATURES = {
    "pebble_last_error": (ctypes.c_char_p, []),
    "pebble_abi_version": (c_int, []),
    "pebble_device_count": (c_int, []),
    "pebble_set_device": (c_int, [c_int]),
    "pebble_device_synchronize": (c_int, []),
    "pebble_launch_count": (c_u64, []),
    # device sample formats (normalizeIQ)
    "pebble_iq_bytes_per_sample": (c_int, [c_int]),
    "pebble_normalize_iq": (c_int, [c_int, c_int, c_dbl, c_dbl, c_int, ctypes.c_longlong, c_vp, c_vp]),
    "pebble_normalize_iq_device": (c_int, [c_int, c_int, c_dbl, c_dbl, c_int, ctypes.c_longlong, c_vp, c_vp, c_vp]),
    "pebble_downconvert_set_input_format": (c_int, [c_vp, c_int, c_int, c_dbl, c_dbl]),
    "pebble_downconvert_input_bytes_per_sample": (c_int, [c_vp]),
    "pebble_downconvert_set_strict_lengths": (c_int, [c_vp, c_int]),
    "pebble_receiver_set_input_format": (c_int, [c_vp, c_int, c_int, c_dbl, c_dbl]),
    # CFastFIR
    "pebble_fastfir_create": (c_vp, []),
    "pebble_fastfir_create_sized": (c_vp, [c_int, c_int]),
    "pebble_fastfir_destroy": (None, [c_vp]),
    "pebble_fastfir_setup": (c_int, [c_vp, c_dbl, c_dbl, c_dbl, c_dbl]),
    "pebble_fastfir_set_fused_scale": (c_int, [c_vp, c_dbl, c_dbl]),
    "pebble_fastfir_process": (c_int, [c_vp, c_int, c_vp, c_vp]),
    "pebble_fastfir_process_device": (c_int, [c_vp, c_int, c_vp, c_vp, c_vp]),
    "pebble_fastfir_out_count": (c_int, [c_vp, c_int]),
    "pebble_fastfir_state_size": (c_i64, [c_vp]),
    "pebble_fastfir_get_state": (c_int, [c_vp, c_vp, c_i64]),
    "pebble_fastfir_set_state": (c_int, [c_vp, c_vp, c_i64]),
    "pebble_fastfir_get_filter": (c_int, [c_vp, c_vp]),
    # CDownConvert
    "pebble_downconvert_create": (c_vp, []),
    "pebble_downconvert_destroy": (None, [c_vp]),
    "pebble_downconvert_set_data_rate": (c_dbl, [c_vp, c_dbl, c_dbl]),
    "pebble_downconvert_set_data_rate_simple": (c_dbl, [c_vp, c_dbl, c_dbl]),
    "pebble_downconvert_set_frequency": (None, [c_vp, c_dbl]),
    "pebble_downconvert_set_cw_offset": (None, [c_vp, c_dbl]),
    "pebble_downconvert_num_stages": (c_int, [c_vp]),
    "pebble_downconvert_plan": (c_int, [c_dbl, c_dbl, c_int, c_vp, c_int, c_vp]),
    "pebble_downconvert_stage_lengths": (c_int, [c_vp, c_vp, c_int]),
    "pebble_downconvert_composite_response": (c_int, [c_dbl, c_dbl, c_int, c_vp, c_int, c_vp]),
    "pebble_downconvert_process": (c_int, [c_vp, c_int, c_vp, c_vp]),
    "pebble_downconvert_process_device": (c_int, [c_vp, c_int, c_vp, c_vp, c_vp]),
    "pebble_downconvert_state_size": (c_i64, [c_vp]),
    "pebble_downconvert_get_state": (c_int, [c_vp, c_vp, c_i64]),
    "pebble_downconvert_set_state": (c_int, [c_vp, c_vp, c_i64]),
    # Decimator
    "pebble_decimator_create": (c_vp, [ctypes.c_uint, ctypes.c_uint]),
    "pebble_decimator_destroy": (None, [c_vp]),
    "pebble_decimator_build_chain": (ctypes.c_float, [c_vp, ctypes.c_uint, ctypes.c_uint, ctypes.c_uint]),
    "pebble_decimator_dec_by2_stages": (c_int, [c_vp]),
    "pebble_decimator_total_decimation": (c_int, [c_vp]),
    "pebble_decimator_chain": (c_int, [c_vp, c_vp, c_vp, c_int]),
    "pebble_decimator_plan": (c_int, [ctypes.c_uint, ctypes.c_uint, ctypes.c_uint, c_vp, c_vp, c_int, c_vp, c_vp]),
    "pebble_decimator_out_count": (c_int, [c_vp, c_int]),
    "pebble_decimator_process": (c_int, [c_vp, c_int, c_vp, c_vp]),
    "pebble_decimator_process_device": (c_int, [c_vp, c_int, c_vp, c_vp, c_vp]),
    "pebble_decimator_state_size": (c_i64, [c_vp]),
    "pebble_decimator_get_state": (c_int, [c_vp, c_vp, c_i64]),
    "pebble_decimator_set_state": (c_int, [c_vp, c_vp, c_i64]),
    # Mixer
    "pebble_mixer_create": (c_vp, [ctypes.c_uint, ctypes.c_uint]),
    "pebble_mixer_destroy": (None, [c_vp]),
    "pebble_mixer_set_frequency": (None, [c_vp, c_dbl]),
    "pebble_mixer_is_passthrough": (c_int, [c_vp]),
    "pebble_mixer_process": (c_int, [c_vp, c_int, c_vp, c_vp]),
    "pebble_mixer_process_device": (c_int, [c_vp, c_int, c_vp, c_vp, c_vp]),
    # NCO + CpxUtil
    "pebble_nco_create": (c_vp, [ctypes.c_uint, ctypes.c_uint]),
    "pebble_nco_destroy": (None, [c_vp]),
    "pebble_nco_set_frequency": (c_dbl, [c_vp, c_dbl]),
    "pebble_nco_frequency": (c_dbl, [c_vp]),
    "pebble_nco_gen_single": (c_int, [c_vp, c_vp, c_int, c_dbl, c_int]),
    "pebble_nco_gen_single_device": (c_int, [c_vp, c_vp, c_int, c_dbl, c_int, c_vp]),
    "pebble_cpx_copy_device": (c_int, [c_vp, c_vp, ctypes.c_longlong, c_vp]),
    "pebble_cpx_scale_device": (c_int, [c_vp, c_vp, c_dbl, ctypes.c_longlong, c_vp]),
    "pebble_cpx_mult_device": (c_int, [c_vp, c_vp, c_vp, ctypes.c_longlong, c_vp]),
    "pebble_cpx_add_device": (c_int, [c_vp, c_vp, c_vp, ctypes.c_longlong, c_vp]),
    "pebble_cpx_mag_device": (c_int, [c_vp, c_vp, ctypes.c_longlong, c_vp]),
    "pebble_cpx_sqrmag_device": (c_int, [c_vp, c_vp, ctypes.c_longlong, c_vp]),
    "pebble_cpx_clear_device": (c_int, [c_vp, ctypes.c_longlong, c_vp]),
    "pebble_cpx_decimate_device": (c_int, [c_vp, c_vp, c_int, ctypes.c_longlong, c_vp]),
    # FFT / SignalSpectrum
    "pebble_fft_create": (c_vp, []),
    "pebble_fft_destroy": (None, [c_vp]),
    "pebble_fft_params": (c_int, [c_vp, ctypes.c_uint, c_dbl, c_dbl, c_int, c_int]),
    "pebble_fft_size": (c_int, [c_vp]),
    "pebble_fft_forward": (c_int, [c_vp, c_vp, c_vp, c_int]),
    "pebble_fft_inverse": (c_int, [c_vp, c_vp, c_vp, c_int]),
    "pebble_fft_spectrum": (c_int, [c_vp, c_vp, c_vp, c_int]),
    "pebble_fft_spectrum_batch_device": (c_int, [c_vp, c_int, c_int, c_vp, c_vp, c_vp, c_vp]),
    "pebble_window": (c_dbl, [c_int, c_int, c_vp]),
    "pebble_fft_get_state": (c_int, [c_vp, c_vp, c_i64]),
    "pebble_fft_set_state": (c_int, [c_vp, c_vp, c_i64]),
    "pebble_fft_set_backend": (c_int, [c_vp, c_int]),
    # SignalSpectrum
    "pebble_spectrum_create": (c_vp, [ctypes.c_uint, ctypes.c_uint, ctypes.c_uint, c_int, c_int, c_int]),
    "pebble_spectrum_destroy": (None, [c_vp]),
    "pebble_spectrum_set_sample_rate": (c_int, [c_vp, ctypes.c_uint, ctypes.c_uint]),
    "pebble_spectrum_set_hires": (None, [c_vp, c_int]),
    "pebble_spectrum_set_updates_per_sec": (None, [c_vp, c_int]),
    "pebble_spectrum_unprocessed": (c_int, [c_vp, c_vp, c_int, ctypes.c_longlong]),
    "pebble_spectrum_zoomed": (c_int, [c_vp, c_vp, c_int, ctypes.c_longlong]),
    "pebble_spectrum_bin_count": (c_int, [c_vp]),
    "pebble_spectrum_hires_bin_count": (c_int, [c_vp]),
    "pebble_spectrum_overload": (c_int, [c_vp]),
    "pebble_spectrum_overrun": (c_int, [c_vp]),
    "pebble_spectrum_ack_update": (None, [c_vp]),
    "pebble_spectrum_get_unprocessed": (c_vp, [c_vp]),
    "pebble_spectrum_get_hires": (c_vp, [c_vp]),
    # CFir
    "pebble_cfir_create": (c_vp, []),
    "pebble_cfir_destroy": (None, [c_vp]),
    "pebble_cfir_init_lp": (c_int, [c_vp, c_int, c_dbl, c_dbl, c_dbl, c_dbl, c_dbl]),
    "pebble_cfir_init_hp": (c_int, [c_vp, c_int, c_dbl, c_dbl, c_dbl, c_dbl, c_dbl]),
    "pebble_cfir_init_const": (c_int, [c_vp, c_int, c_vp, c_vp, c_dbl]),
    "pebble_cfir_get_taps": (c_int, [c_vp, c_vp, c_int]),
    "pebble_cfir_process": (c_int, [c_vp, c_int, c_vp, c_vp]),
    "pebble_cfir_process_device": (c_int, [c_vp, c_int, c_vp, c_vp, c_vp]),
    # AGC
    "pebble_agc_create": (c_vp, [ctypes.c_uint, ctypes.c_uint]),
    "pebble_agc_destroy": (None, [c_vp]),
    "pebble_agc_set_mode": (c_int, [c_vp, c_int, c_int]),
    "pebble_agc_set_threshold": (c_int, [c_vp, c_int]),
    "pebble_agc_process": (c_int, [c_vp, c_int, c_vp, c_vp]),
    "pebble_agc_process_device": (c_int, [c_vp, c_int, c_vp, c_vp, c_vp]),
    # Demod
    "pebble_demod_create": (c_vp, [ctypes.c_uint, ctypes.c_uint]),
    "pebble_demod_destroy": (None, [c_vp]),
    "pebble_demod_set_mode": (c_int, [c_vp, c_int]),
    "pebble_demod_set_nfm_variant": (c_int, [c_vp, c_int]),
    "pebble_demod_set_wfm_sample_rate": (c_int, [c_vp, c_dbl]),
    "pebble_demod_get_stereo_lock": (c_int, [c_vp]),
    "pebble_demod_set_rds": (c_int, [c_vp, c_int]),
    "pebble_demod_get_rds_rate": (c_dbl, [c_vp]),
    "pebble_demod_get_next_rds_group": (c_int, [c_vp, c_vp]),
    "pebble_demod_get_rds_tap": (c_int, [c_vp, c_vp, c_int]),
    "pebble_rds_sync_create": (c_vp, []),
    "pebble_rds_sync_destroy": (None, [c_vp]),
    "pebble_rds_sync_push_bits": (c_int, [c_vp, c_vp, c_int]),
    "pebble_rds_sync_next_group": (c_int, [c_vp, c_vp]),
    "pebble_demod_set_bandwidth": (c_int, [c_vp, c_dbl]),
    "pebble_demod_is_passthrough": (c_int, [c_vp]),
    "pebble_demod_process": (c_int, [c_vp, c_int, c_vp, c_vp]),
    "pebble_demod_process_device": (c_int, [c_vp, c_int, c_vp, c_vp, c_vp]),
    # batched forms + channelizer
    "pebble_downconvert_set_channels": (c_int, [c_vp, c_int]),
    "pebble_downconvert_set_channel_frequency": (c_int, [c_vp, c_int, c_dbl]),
    "pebble_downconvert_process_channels_device": (c_int, [c_vp, c_int, c_vp, c_vp, ctypes.c_longlong, c_vp]),
    "pebble_downconvert_batch_path": (c_int, [c_vp]),
    "pebble_fastfir_set_streams": (c_int, [c_vp, c_int]),
    "pebble_fastfir_process_batch_device": (c_int, [c_vp, c_int, c_vp, ctypes.c_longlong, c_vp, ctypes.c_longlong, c_vp]),
    "pebble_channelizer_create": (c_vp, [c_int, c_dbl, c_dbl]),
    "pebble_channelizer_destroy": (None, [c_vp]),
    "pebble_channelizer_out_rate": (c_dbl, [c_vp]),
    "pebble_channelizer_num_stages": (c_int, [c_vp]),
    "pebble_channelizer_num_channels": (c_int, [c_vp]),
    "pebble_channelizer_set_frequency": (c_int, [c_vp, c_int, c_dbl]),
    "pebble_channelizer_set_filter": (c_int, [c_vp, c_dbl, c_dbl]),
    "pebble_channelizer_set_demod_mode": (c_int, [c_vp, c_int, c_int]),
    "pebble_channelizer_set_am_bandwidth": (c_int, [c_vp, c_dbl]),
    "pebble_channelizer_set_agc_threshold": (c_int, [c_vp, c_int]),
    "pebble_channelizer_out_count": (c_int, [c_vp, c_int]),
    "pebble_channelizer_set_squelch": (c_int, [c_vp, c_vp, c_int]),
    "pebble_channelizer_squelch_from_spectrum": (c_int, [c_vp, c_vp, c_int, ctypes.c_uint, c_dbl, c_vp]),
    "pebble_channelizer_set_input_format": (c_int, [c_vp, c_int, c_int, c_dbl, c_dbl]),
    "pebble_channelizer_front_path": (c_int, [c_vp]),
    "pebble_channelizer_downconvert_device": (c_int, [c_vp, c_int, c_vp, c_vp, ctypes.c_longlong, c_vp]),
    "pebble_channelizer_process_device": (c_int, [c_vp, c_int, c_vp, c_vp, ctypes.c_longlong, c_vp]),
    # CFractResampler
    "pebble_resampler_create": (c_vp, [c_int]),
    "pebble_resampler_destroy": (None, [c_vp]),
    "pebble_resampler_set_streams": (c_int, [c_vp, c_int]),
    "pebble_resampler_out_count": (c_int, [c_vp, c_int, c_int, c_dbl]),
    "pebble_resampler_resample": (c_int, [c_vp, c_int, c_dbl, c_vp, c_vp]),
    "pebble_resampler_resample_device": (c_int, [c_vp, c_int, c_int, c_dbl, c_vp, ctypes.c_longlong, c_vp, ctypes.c_longlong, c_int, c_vp]),
    "pebble_resampler_resample_stereo16_device": (c_int, [c_vp, c_int, c_int, c_dbl, c_vp, ctypes.c_longlong, c_vp, ctypes.c_longlong, c_dbl, c_vp]),
    "pebble_resampler_state_size": (c_i64, [c_vp]),
    "pebble_resampler_get_state": (c_int, [c_vp, c_vp, c_i64]),
    "pebble_resampler_set_state": (c_int, [c_vp, c_vp, c_i64]),
    "pebble_resampler_plan": (c_int, [c_dbl, c_dbl, c_int, c_vp]),
    "pebble_resampler_plan_times": (c_int, [c_dbl, c_dbl, c_int, c_vp, c_int]),
    # CIir / IQBalance / squelch
    "pebble_iir_create": (c_vp, []),
    "pebble_iir_destroy": (None, [c_vp]),
    "pebble_iir_init": (c_int, [c_vp, c_int, c_dbl, c_dbl, c_dbl]),
    "pebble_iir_get_coefficients": (c_int, [c_vp, c_vp]),
    "pebble_iir_set_coefficients": (c_int, [c_vp, c_vp]),
    "pebble_iir_process": (c_int, [c_vp, c_int, c_vp, c_vp]),
    "pebble_iir_process_device": (c_int, [c_vp, c_int, c_vp, c_vp, c_vp]),
    "pebble_iir_get_state": (c_int, [c_vp, c_vp]),
    "pebble_iir_set_state": (c_int, [c_vp, c_vp]),
    "pebble_iqbalance_process_device": (c_int, [c_int, c_int, c_dbl, c_dbl, c_vp, c_vp, c_vp]),
    "pebble_sigstrength_fd_estimate_device": (c_int, [c_vp, c_int, ctypes.c_uint, ctypes.c_float, ctypes.c_float, c_vp, c_int, c_dbl, c_vp, c_vp]),
    "pebble_sigstrength_fd_estimate_noalloc_max_channels": (c_int, []),
    "pebble_sigstrength_fd_estimate_noalloc_device": (c_int, [c_vp, c_int, ctypes.c_uint, ctypes.c_float, ctypes.c_float, c_vp, c_int, c_dbl, c_vp, c_vp]),
    # NoiseBlanker
    "pebble_noiseblanker_create": (c_vp, [ctypes.c_uint, ctypes.c_uint]),
    "pebble_noiseblanker_destroy": (None, [c_vp]),
    "pebble_noiseblanker_set_nb_enabled": (c_int, [c_vp, c_int]),
    "pebble_noiseblanker_set_nb2_enabled": (c_int, [c_vp, c_int]),
    "pebble_noiseblanker_enabled": (c_int, [c_vp, c_int]),
    "pebble_noiseblanker_process": (c_int, [c_vp, c_int, c_int, c_vp, c_vp]),
    "pebble_noiseblanker_process_device": (c_int, [c_vp, c_int, c_int, c_vp, c_vp, c_vp]),
    # Receiver chain
    "pebble_receiver_create": (c_vp, [c_dbl, c_int]),
    "pebble_receiver_destroy": (None, [c_vp]),
    "pebble_receiver_set_data_rate": (c_dbl, [c_vp, c_dbl]),
    "pebble_receiver_set_mixer": (None, [c_vp, c_dbl]),
    "pebble_receiver_set_filter": (c_int, [c_vp, c_dbl, c_dbl]),
    "pebble_receiver_set_demod_mode": (c_int, [c_vp, c_int]),
    "pebble_receiver_set_demod_bandwidth": (c_int, [c_vp, c_dbl]),
    "pebble_receiver_set_agc": (c_int, [c_vp, c_int, c_int]),
    "pebble_receiver_set_gain_stages": (c_int, [c_vp, c_int]),
    "pebble_receiver_out_count": (c_int, [c_vp, c_int]),
    "pebble_receiver_set_audio": (c_int, [c_vp, c_int, c_int]),
    "pebble_receiver_get_wfm_rate": (c_dbl, [c_vp]),
    "pebble_receiver_demod": (c_vp, [c_vp]),
    "pebble_receiver_set_profiling": (c_int, [c_vp, c_int]),
    "pebble_receiver_get_profile": (c_int, [c_vp, c_vp, c_vp]),
    "pebble_receiver_process": (c_int, [c_vp, c_int, c_vp, c_vp]),
    "pebble_receiver_process_device": (c_int, [c_vp, c_int, c_vp, c_vp, c_vp]),
    "pebble_receiver_process_pinned": (c_int, [c_vp, c_int, c_vp, c_vp, c_vp]),
}


def load(path=LIB_PATH):
    if not os.path.exists(path):
        raise ImportError(
            "pebblesdr_b200: %s is missing. Build it with `python pebblesdr_b200/build.py` "
            "(nvcc, sm_100a). There is no CPU fallback." % path)
    lib = ctypes.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError here == header/library mismatch: fail loudly
        fn.restype = res
        fn.argtypes = args
    return lib


lib = load()


class PebbleError(RuntimeError):
    pass


def check(rc, what=""):
    """Negative return codes are errors (PEBBLE_ERR_*)."""
    if rc is None or (isinstance(rc, int) and rc < 0):
        raise PebbleError("%s failed (%s): %s" % (what, rc, lib.pebble_last_error().decode()))
    return rc
