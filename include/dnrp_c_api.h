/* dnrp_c_api.h -- C ABI of libdnrp_cuda.so, the B200 (sm_100a) implementation of the DECT NR+ PCC/PDC
 * physical-layer hot path.  Plain pointers and sizes only; every function returns 0 on success or a
 * negative dnrp_status_t; no exceptions cross this boundary; there is NO CPU fallback behind it (a
 * missing GPU or kernel image is an error).
 *
 * The reference (maxpenner/DECT-NR-Plus-SDR) has no FFI layer: firmware is compiled in (reference
 * README.md:217).  Each entry point below therefore names the reference C++ interface it replaces
 * (file:line relative to /root/reference); INTEGRATION.md shows the binding a maintainer would add.
 *
 * Batching model: one context = one device + one packet geometry ("plan": psdef, antennas, resampling)
 * + a maximum batch size (<= 65535: packets ride in a CUDA grid dimension).  A call processes n <= max_batch packets
 * of that geometry; network id and PLCF type may differ per packet (the *_ex calls); packets of several geometries
 * are bucketed per context by the host adapter (dnrp_host.hpp, batch_phy_t).  Buffers passed in
 * are HOST pointers unless the name ends in _dev.
 */
#ifndef DNRP_C_API_H
#define DNRP_C_API_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    DNRP_OK = 0,
    DNRP_ERR_INVALID = -1,   /* invalid argument / impossible packet configuration */
    DNRP_ERR_CUDA = -2,      /* CUDA runtime error, see dnrp_last_error_string() */
    DNRP_ERR_NO_DEVICE = -3, /* no usable sm_100 device */
    DNRP_ERR_STATE = -4      /* call order violated (e.g. PDC before PCC) */
} dnrp_status_t;

typedef struct dnrp_ctx dnrp_ctx_t;
typedef struct { float re, im; } dnrp_cf32_t;

/* sp3::packet_sizes_def_t, lib/include/dectnrp/sections_part3/derivative/packet_sizes_def.hpp */
typedef struct {
    uint32_t u, b, PacketLengthType, PacketLength, tm_mode_index, mcs_index, Z;
} dnrp_psdef_t;

/* subset of sp3::packet_sizes_t, lib/include/dectnrp/sections_part3/derivative/packet_sizes.hpp */
typedef struct {
    dnrp_psdef_t psdef;
    uint32_t N_eff_TX, N_SS, N_TS, N_TX, closed_loop, N_bps;
    uint32_t N_b_DFT, N_b_CP, N_b_OCC;
    uint32_t N_PACKET_symb, N_DF_symb, N_PDC_subc, N_DRS_subc;
    uint32_t G, N_PDC_bits, N_TB_bits, N_TB_byte, C, K_plus, K_minus, C_plus, C_minus;
    uint32_t N_samples_STF, N_samples_STF_CP_only, N_samples_DF, N_samples_GI;
    uint32_t N_samples_packet_no_GI, N_samples_packet;
} dnrp_packet_sizes_t;

/* replaces worker_pool_config_t (lib/include/dectnrp/phy/pool/worker_pool_config.hpp:35-143) for the
 * parts the hot path needs */
typedef struct {
    dnrp_psdef_t psdef;
    uint32_t N_RX;             /* receive antennas */
    uint32_t oversampling;     /* os_min: 1, 2, 4 ... (N_b_DFT_os = oversampling * N_b_DFT) */
    uint32_t resample_L, resample_M; /* DECT rate * oversampling * L / M = hardware rate (1/1 = no resampling); with
                                      * L != M every IQ buffer crossing this ABI is at the hardware rate: TX resamples
                                      * up (tx.cpp:913-934), sync and the receiver resample down (rx_pacer.cpp) */
    uint32_t max_batch;
    uint32_t network_id;       /* worker_pool_t::add_network_id, worker_pool.cpp:109-114 */
    uint32_t plcf_type;        /* 1 or 2: PLCF type transmitted / PDC scrambling variant */
    uint32_t chestim_mode_lr;  /* phy.json chestim_mode_lr_default */
    uint32_t chestim_mode_lr_t_stride; /* phy.json chestim_mode_lr_t_stride_default */
    uint32_t tdec_min_iter, tdec_max_iter; /* 0 = reference defaults 2 / 10 (pdc_enc.cpp:35-36) */
    /* Limited soft buffer ("channel coding with a limited number of soft bits is not implemented yet", reference
     * README.md:180; ToDo at lib/src/phy/fec/pdc_enc.cpp:122-123 pointing to ETSI TS 103 636-3 6.1.5.3, which follows
     * TS 36.212 5.1.4.1.2).  N_soft = soft bits of the receiving radio device (radio_device_class_t::N_soft_min,
     * radio_device_class.cpp:37-141, or the RD capability IE), M_DL_HARQ = HARQ processes sharing them: a transport
     * block may use N_IR = N_soft / M_DL_HARQ soft bits, each of its C code blocks a circular buffer of
     * N_cb = min(floor(N_IR / C), K_w); transmitter (bit selection) and receiver (soft combining) both honour it, and
     * every redundancy version starts at k0(N_cb).  N_soft = 0: no limitation (N_cb = K_w), the reference's behaviour.
     * Own activation of a planned feature: DESIGN.md 4b. */
    uint32_t N_soft, M_DL_HARQ;
} dnrp_config_t;

/* sp3::get_packet_sizes, lib/src/sections_part3/derivative/packet_sizes.cpp:97-227.
 * DNRP_ERR_INVALID where the reference returns std::nullopt. Host only. */
int dnrp_get_packet_sizes(const dnrp_psdef_t* psdef, dnrp_packet_sizes_t* out);

/* resampler_t's filter design (lib/src/phy/resample/resampler.cpp:55-140, filter/kaiser.cpp:37-118,
 * resampler_param.hpp:61-86) for L/M at the given oversampling.  Host only.  taps_out [subfilter_length][L] =
 * L * h[l + j L], zero padded (may be NULL); info4 = {filter_length, subfilter_length, delay, taps written}.
 * dnrp_resampler_n_after = resampler_t::get_N_samples_after_resampling (resampler.cpp:310-328). */
int dnrp_resampler_design(uint32_t L, uint32_t M, uint32_t oversampling, float* taps_out, uint32_t taps_cap,
                          uint32_t* info4);
uint32_t dnrp_resampler_n_after(uint32_t L, uint32_t M, uint32_t oversampling, uint32_t n_in);

/* tx_rx_t / worker_tx_rx_t construction (lib/src/phy/tx_rx.cpp:38-129, pool/worker_tx_rx.cpp:43-67):
 * builds all LUTs and device workspaces for one geometry */
int dnrp_ctx_create(const dnrp_config_t* cfg, int device, dnrp_ctx_t** out);
void dnrp_ctx_destroy(dnrp_ctx_t* ctx);
const char* dnrp_last_error_string(void);
int dnrp_ctx_get_packet_sizes(const dnrp_ctx_t* ctx, dnrp_packet_sizes_t* out);
int dnrp_ctx_synchronize(dnrp_ctx_t* ctx);

/* ---- FEC only (fec_t, lib/include/dectnrp/phy/fec/fec.hpp:84-166) --------------------------------
 * encode_plcf + encode_tb for n packets: plcf [n][10] bytes (type 1 uses the first 5), cl_bf_mask [n]
 * CRC masks (0, 0x5555, 0xAAAA, 0xFFFF; NULL = 0), tb [n][N_TB_byte], rv [n] (NULL = 0).
 * Outputs: pcc_bits [n][28] (196 bits used), pdc_bits [n][ceil(G/8) rounded up to 4]. */
int dnrp_fec_encode(dnrp_ctx_t* ctx, uint32_t n, const uint8_t* plcf, const uint32_t* cl_bf_mask,
                    const uint8_t* tb, const uint8_t* rv, uint8_t* pcc_bits, uint8_t* pdc_bits);

/* decode_tb for n packets from int16 LLRs [n][G] (LLR > 0 <=> bit 1, reference tb2pdc.cpp:173-176).
 * harq_slot [n] (NULL: slot = packet index), rv [n] (NULL = 0), harq_reset [n] (NULL = 1: first
 * transmission).  Outputs: tb_out [n][N_TB_byte], crc_ok [n], n_iter [n][C] (may be NULL). */
int dnrp_fec_decode_tb(dnrp_ctx_t* ctx, uint32_t n, const int16_t* llr, const uint32_t* harq_slot,
                       const uint8_t* rv, const uint8_t* harq_reset, uint8_t* tb_out, uint8_t* crc_ok,
                       uint8_t* n_iter);

/* decode_plcf_test for both PLCF types from 196 LLRs per packet (pcc_enc.cpp:215-364).
 * Outputs per packet and type t = 0 (type 1), 1 (type 2): ok [n][2], plcf [n][2][10],
 * mask [n][2] (matched CRC mask), n_iter [n][2] (may be NULL). */
int dnrp_fec_decode_plcf(dnrp_ctx_t* ctx, uint32_t n, const int16_t* llr196, uint8_t* ok, uint8_t* plcf,
                         uint32_t* mask, uint8_t* n_iter);

/* dnrp_fec_decode_plcf for callers that must not allocate: code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc.  NO COUNTERPART IN THE REFERENCE.
 * Same kernels and results as dnrp_fec_decode_plcf, but (i) inputs and outputs are DEVICE pointers owned by the
 * caller, (ii) the decoder's scratch (the de-rate-matched soft buffers, 1280 bytes per packet) lives in a static
 * __device__ array of DNRP_NOALLOC_MAX_PACKETS packets (5 MiB, held by every GPU that loads the library until the
 * process exits) instead of a context workspace, (iii) the call only enqueues two kernels on `stream` (a
 * cudaStream_t; NULL = the context's stream) and returns without synchronising; it calls no cudaMalloc / cudaFree /
 * allocating library routine.  n > DNRP_NOALLOC_MAX_PACKETS returns DNRP_ERR_INVALID.  All calls on one GPU share
 * the scratch array (it keeps the last call's contents; every value a decode reads is rewritten first, nothing needs
 * clearing): calls that overlap on the GPU (different streams, or graphs replayed concurrently) RACE on it --
 * serialise them.  The first call per device looks the array's device address up (cudaGetSymbolAddress); make that
 * call outside stream capture.
 * llr196_dev [n][196] int16; ok_dev [n][2]; plcf_dev [n][2][16] (the first 5 / 10 bytes of each entry are the PLCF);
 * mask_dev [n][2] (matched CRC mask, 0xffffffff = none); n_iter_dev [n][2]. */
#define DNRP_NOALLOC_MAX_PACKETS 4096
int dnrp_fec_decode_plcf_noalloc(dnrp_ctx_t* ctx, uint32_t n, const int16_t* llr196_dev, uint8_t* ok_dev,
                                 uint8_t* plcf_dev, uint32_t* mask_dev, uint8_t* n_iter_dev, void* stream);

/* ---- transmitter ------------------------------------------------------------------------------
 * dnrp_tx_desc_t = tx_descriptor_t{codebook_index, tx_meta_t} + the HARQ redundancy version
 * (lib/include/dectnrp/phy/tx/tx_descriptor.hpp:43-78, tx_meta.hpp:32-58, harq/process_tx.hpp). */
typedef struct {
    uint32_t codebook_index;
    uint32_t optimal_scaling_DAC;
    float DAC_scale;
    float iq_phase_rad;
    float iq_phase_increment_s2s_post_resampling_rad;
    uint32_t GI_percentage;
    uint32_t rv;
} dnrp_tx_desc_t;

/* samples per antenna reserved for one packet in iq_out (packet incl. full guard interval, padded) */
uint32_t dnrp_tx_stride(const dnrp_ctx_t* ctx);

/* tx_t::generate_tx_packet (lib/src/phy/tx/tx.cpp:165-314) for n packets.
 * plcf [n][10], tb [n][N_TB_byte]; iq_out [n][N_TX][dnrp_tx_stride()] (zero padded; may be NULL to keep
 * the IQ on the device for dnrp_channel_batch); n_samples_out [n] = N_samples_transmit (may be NULL). */
int dnrp_tx_batch(dnrp_ctx_t* ctx, uint32_t n, const dnrp_tx_desc_t* descs, const uint8_t* plcf,
                  const uint8_t* tb, dnrp_cf32_t* iq_out, uint32_t* n_samples_out);

/* dnrp_tx_batch with the IQ in the radio's wire format: iq_out [n][N_TX][dnrp_tx_stride()][2] int16 (I, Q) =
 * round-to-nearest(sample * scale), saturated (scale = 32767 corresponds to UHD's fc32 -> sc16 conversion, which the
 * reference leaves to UHD on the host CPU: lib/src/radio/hw_usrp.cpp:422-424, cpu_format "fc32", otw_format "sc16").
 * Halves the device-to-host bytes; the float IQ stays on the device as after dnrp_tx_batch.  dnrp_tx_stride() must be a
 * multiple of 4. */
int dnrp_tx_batch_sc16(dnrp_ctx_t* ctx, uint32_t n, const dnrp_tx_desc_t* descs, const uint8_t* plcf,
                       const uint8_t* tb, int16_t* iq_out, float scale, uint32_t* n_samples_out);

/* ---- synchronised receiver ----------------------------------------------------------------------
 * dnrp_sync_report_t = the fields of sync_report_t the synchronised receiver consumes
 * (lib/include/dectnrp/phy/rx/sync/sync_report.hpp:41-111).  fine_peak_time_local is the index of the
 * first STF sample inside this packet's IQ window (the reference's fine_peak_time_64 relative to the
 * buffer position the window starts at); the cfo values are CORRECTION phase increments in rad per
 * sample at the oversampled rate, as sync_t hands them over; rms[a] = 0 lets the STF stage estimate it. */
typedef struct {
    uint32_t fine_peak_time_local;
    float cfo_fractional_rad;
    float cfo_integer_rad;
    float rms[8];
} dnrp_sync_report_t;

/* pcc_report_t (lib/include/dectnrp/phy/rx/rx_synced/pcc_report.hpp:33-52) + the sync_report_t fields
 * the PCC stage updates (rx_synced.cpp:258-262): index 0 = PLCF type 1, 1 = type 2 */
typedef struct {
    uint8_t plcf_ok[2];
    uint8_t plcf[2][10];
    uint32_t crc_mask[2];       /* matched mask: 0 none, 0x5555 closed loop, 0xAAAA beamforming, 0xFFFF both */
    float snr_dB;
    float cfo_residual_rad;     /* STF based residual, added to cfo_fractional_rad from then on */
    float sto_fractional;       /* in samples */
    float rms[8];
} dnrp_pcc_report_t;

/* pdc_report_t (lib/include/dectnrp/phy/rx/rx_synced/pdc_report.hpp:33-56) */
typedef struct {
    uint32_t crc_ok;
    float snr_dB;
    /* mimo_report_t (rx_synced/mimo/mimo_report.hpp:27-59, estimator_mimo.cpp:80-272): codebook indices recommended
     * for single-stream beamforming (modes 3 / 7), searched on the last DRS cells of the packet; 0 where no
     * beamforming codebook exists.  The reference fills it for CRC-passing packets only; here always. */
    uint32_t mimo_N_RX, mimo_N_TS_other;
    uint32_t tm_3_7_beamforming_idx, tm_3_7_beamforming_reciprocal_idx;
} dnrp_pdc_report_t;

/* samples per antenna reserved for one packet's IQ window in iq_in */
uint32_t dnrp_rx_stride(const dnrp_ctx_t* ctx);

/* ---- packet synchronisation -------------------------------------------------------------------
 * sync_report_t as sync_chunk_t::search fills it (lib/src/phy/rx/sync/sync_chunk.cpp:120-215,
 * lib/include/dectnrp/phy/rx/sync/sync_report.hpp:41-111), times relative to the window start. */
typedef struct {
    uint32_t detected;                 /* 0: no packet in this window, the other fields are zero */
    uint32_t detection_time_local;
    uint32_t detection_ant_idx;
    float detection_metric;
    float detection_rms;
    uint32_t coarse_peak_time_local;
    float coarse_peak_array[8];        /* 0 for antennas whose coarse peak was not valid */
    float rms_array[8];
    float cfo_fractional_rad;
    float cfo_integer_rad;
    uint32_t N_eff_TX;
    uint32_t fine_peak_time_local;
    uint32_t n_detections;             /* detections in this window incl. the ones dropped as invalid */
    uint32_t strongest_ant;            /* antenna the cross-correlation ran on */
    uint32_t coarse_peak_time_hw;      /* coarse peak converted to the hardware rate (= local without resampling) */
} dnrp_sync_result_t;

/* sync_chunk_t::search for n IQ windows, first packet of each window: autocorrelator_detection_t,
 * autocorrelator_peak_t (coarse peak, fractional CFO, rms), crosscorrelator_t (N_eff_TX, fine peak); constants of
 * sync_param.hpp as shipped.  iq_in [n][N_RX][dnrp_rx_stride()] host windows or NULL (windows already on the
 * device).  search_length = samples the detection may consume (0: window minus one STF and one pattern).
 * out [n] may be NULL.  The results also stay on the device: a following dnrp_rx_pcc_batch with sync = NULL
 * runs on them (packets not detected are processed at offset 0 and fail their CRCs). */
int dnrp_sync_batch(dnrp_ctx_t* ctx, uint32_t n, const dnrp_cf32_t* iq_in, uint32_t search_length,
                    dnrp_sync_result_t* out);

/* dnrp_sync_batch for host windows in the radio's wire format, [n][N_RX][dnrp_rx_stride()][2] int16 (I, Q), converted on
 * the device (sample = (I, Q) * scale; 1/32768 reproduces UHD's sc16 -> fc32 conversion, hw_usrp.cpp:436-438): half the
 * host-link bytes for a caller that keeps the reference's flow sync -> PCC -> work_pcc -> PDC (dnrp_rx_pcc_batch and
 * dnrp_rx_pdc_batch continue on the converted windows with iq_in = NULL).  dnrp_rx_stride() must be a multiple of 4. */
int dnrp_sync_batch_sc16(dnrp_ctx_t* ctx, uint32_t n, const int16_t* iq_in, float scale, uint32_t search_length,
                         dnrp_sync_result_t* out);


/* Streaming front door: packet search over ONE continuous IQ stream instead of pre-cut windows -- the job of
 * worker_sync_t / sync_chunk_t / baton_t (lib/src/phy/pool/worker_sync.cpp:112-221, src/phy/rx/sync/sync_chunk.cpp:
 * 146-283, src/phy/pool/baton.cpp:157-169).  The stream is cut into chunks of chunk_hop samples (0: four packet
 * lengths) that overlap by one STF + one pattern of history in front and the peak / cross-correlation range behind;
 * all chunks are searched in parallel; a chunk that found a packet continues behind it (skip_after_peak: 2 STFs) in
 * the next round until its detection range is used up; results are ordered by time and double detections closer than
 * one STF pattern to their predecessor are dropped (RX_SYNC_PARAM_SYNC_TIME_UNIQUE_LIMIT_IN_STF_PATTERNS_DP).
 * iq [N_RX][n_samples] host samples at the hardware rate (contexts without resampling only; dnrp_stream_push below
 * handles any L / M).  A chunk_hop whose detection prefix sums exceed the device's shared memory is refused with
 * DNRP_ERR_INVALID.  The stream stays on
 * the device for dnrp_stream_extract.  out [max_out]: times are sample indices of the stream. */
typedef struct {
    int64_t fine_peak_time_64;         /* first sample of the packet (STF start) */
    int64_t coarse_peak_time_64;
    dnrp_sync_result_t sync;           /* chunk-local record of the search that found it */
    uint32_t chunk, round;
} dnrp_stream_packet_t;
int dnrp_stream_sync(dnrp_ctx_t* ctx, uint64_t n_samples, const dnrp_cf32_t* iq, uint32_t chunk_hop, uint32_t max_out,
                     dnrp_stream_packet_t* out, uint32_t* n_out);

/* The same search with state, for a radio that delivers its samples piece by piece (the reference's RX ring buffer,
 * radio/buffer_rx.hpp, read by worker_sync_t chunk after chunk; resampling with carried history, rx_pacer.cpp:106-143,
 * 265-327).  dnrp_stream_open sizes a ring on the device for pushes of up to max_push_samples hardware-rate samples per
 * antenna (chunk_hop as above, in DECT-rate samples) and resets the stream clock; dnrp_stream_push appends
 * iq [N_RX][n_samples], resamples the new part to the DECT rate when resample_L != resample_M (absolute phase: the
 * result equals resampling the whole stream at once), searches every detection position whose coarse / fine peak
 * range has fully arrived, and keeps for the next call: the detector's history, the unsearched tail, and the samples
 * of packets whose end has not arrived.  out [max_out] receives, in time order, the packets that are COMPLETE (last
 * sample of the packet including its guard interval is in the ring): times are hardware-rate sample indices counted
 * from the first sample ever pushed; a packet found in one push may be handed out by a later one; complete packets
 * beyond max_out stay queued (push with n_samples = 0 to drain).  dnrp_stream_extract reads the ring, so call it for
 * the packets of one push before the next push.  Results do not depend on how the stream is cut into pushes, except
 * for double detections within one STF pattern, which are dropped in either case. */
int dnrp_stream_open(dnrp_ctx_t* ctx, uint32_t max_push_samples, uint32_t chunk_hop);
/* dnrp_stream_push_sc16: the same for samples in the radio's wire format, iq [N_RX][n_samples][2] int16 (I, Q), sample =
 * (I, Q) * scale on the device (1/32768 = UHD's sc16 -> fc32, hw_usrp.cpp:436-438): half the host-link bytes */
int dnrp_stream_push_sc16(dnrp_ctx_t* ctx, uint32_t n_samples, const int16_t* iq, float scale, uint32_t max_out,
                          dnrp_stream_packet_t* out, uint32_t* n_out);
int dnrp_stream_push(dnrp_ctx_t* ctx, uint32_t n_samples, const dnrp_cf32_t* iq, uint32_t max_out, dnrp_stream_packet_t* out,
                     uint32_t* n_out);
int dnrp_stream_close(dnrp_ctx_t* ctx);

/* rx_pacer_t hand-over to rx_synced_t for n packets found by dnrp_stream_sync / dnrp_stream_push: copies, per packet i, the samples
 * [start_64[i] - t0, start_64[i] - t0 + dnrp_rx_stride()) of the device-resident stream into IQ window i (zeros
 * outside the stream).  dnrp_rx_pcc_batch(ctx, n, NULL, sync, ...) with sync[i].fine_peak_time_local = t0 continues. */
int dnrp_stream_extract(dnrp_ctx_t* ctx, uint32_t n, const int64_t* start_64, uint32_t t0);

/* rx_synced_t::demoddecod_rx_pcc (lib/src/phy/rx/rx_synced/rx_synced.cpp:115-274) for n packets.
 * iq_in [n][N_RX][dnrp_rx_stride()] host IQ windows (NULL: the windows already resident on the device,
 * e.g. written by dnrp_channel_batch); sync [n] or NULL (use the results of the preceding dnrp_sync_batch,
 * which stay on the device); out [n].  Runs the whole front end of the packet (STF
 * stage, all data field FFTs, DRS zero forcing, SNR classes) and the blind PLCF decoding. */
int dnrp_rx_pcc_batch(dnrp_ctx_t* ctx, uint32_t n, const dnrp_cf32_t* iq_in, const dnrp_sync_report_t* sync,
                      dnrp_pcc_report_t* out);

/* rx_synced_t::demoddecod_rx_pdc (rx_synced.cpp:276-361) for the n packets of the preceding
 * dnrp_rx_pcc_batch call: channel interpolation, MRC / Alamouti combining, demapping, decode_tb.
 * harq_slot / rv / harq_reset as in dnrp_fec_decode_tb; tb_out [n][N_TB_byte]; n_iter [n][C] or NULL. */
int dnrp_rx_pdc_batch(dnrp_ctx_t* ctx, uint32_t n, const uint32_t* harq_slot, const uint8_t* rv,
                      const uint8_t* harq_reset, uint8_t* tb_out, dnrp_pdc_report_t* out, uint8_t* n_iter);

/* ---- several networks / PLCF types per batch ---------------------------------------------------------------
 * worker_pool_t::add_network_id -> tx_rx_t::add_new_network_id (lib/src/phy/worker_pool.cpp:109-114, tx_rx.cpp):
 * precomputes the PDC scrambling sequences of both PLCF types of this id (scrambling_pdc.cpp:43-56: c_init = id & 0xff
 * for type 1, id >> 8 for type 2).  The context's own network_id is registered at creation.  Not on the hot path. */
int dnrp_add_network_id(dnrp_ctx_t* ctx, uint32_t network_id);

/* dnrp_tx_batch with the network id and PLCF type taken PER PACKET, as tx_t reads them from each packet's HARQ process
 * (tx.cpp:176-186: hp_tx.get_PLCF_type(), get_network_id()).  network_id [n] / plcf_type [n] (1 or 2); NULL = the
 * context's value.  DNRP_ERR_STATE if an id was not registered with dnrp_add_network_id. */
int dnrp_tx_batch_ex(dnrp_ctx_t* ctx, uint32_t n, const dnrp_tx_desc_t* descs, const uint32_t* network_id,
                     const uint8_t* plcf_type, const uint8_t* plcf, const uint8_t* tb, dnrp_cf32_t* iq_out,
                     uint32_t* n_samples_out);

/* dnrp_rx_pdc_batch with network id and PLCF type per packet, as rx_synced_t::demoddecod_rx_pdc reads them from
 * maclow_phy.hp_rx (rx_synced.cpp:325-350): the firmware's work_pcc decides both from the decoded PLCF. */
int dnrp_rx_pdc_batch_ex(dnrp_ctx_t* ctx, uint32_t n, const uint32_t* harq_slot, const uint8_t* rv,
                         const uint8_t* harq_reset, const uint32_t* network_id, const uint8_t* plcf_type,
                         uint8_t* tb_out, dnrp_pdc_report_t* out, uint8_t* n_iter);

/* Several packet geometries on one receiver.  The reference processes the PCC under the pool's MAXIMUM packet sizes
 * and learns the packet's own geometry from the decoded PLCF (tpoint.cpp:119-141, worksub_psdef); rx_synced_t then
 * takes the packet sizes from hp_rx (rx_synced.cpp:325-350).  Here a context has one geometry: the host runs sync + PCC
 * on the context of the longest packet, then copies the IQ windows of the packets that announced another geometry into
 * the context created for it -- on the device, window src_idx[i] of `src` becomes window i of `dst` (zero filled
 * behind the shorter window) -- and runs dnrp_rx_pcc_batch (with the sync reports of the first pass) +
 * dnrp_rx_pdc_batch there.  Both contexts: same device, mu, beta, oversampling, resampling, N_RX.
 * dnrp_host.hpp (batch_phy_t::add_plan) does all of this behind the tpoint_t callbacks. */
int dnrp_rx_import_windows(dnrp_ctx_t* dst, dnrp_ctx_t* src, uint32_t n, const uint32_t* src_idx);

/* process_rx_t::reset_and_terminate_func (lib/src/phy/harq/process_rx.cpp:56-66) for one HARQ slot: soft values, code
 * block CRC flags and saved code blocks are cleared (a decode with harq_reset = 1 does the same implicitly). */
int dnrp_harq_reset(dnrp_ctx_t* ctx, uint32_t harq_slot);

/* The whole receive chain for n host IQ windows in ONE call: dnrp_sync_batch -> dnrp_rx_pcc_batch ->
 * dnrp_rx_pdc_batch for every window (no MAC decision between PCC and PDC: the simulator's / a monitor's use, where
 * every PDC is wanted; firmware that decides per packet uses the three calls, see dnrp_host.hpp).  The windows are
 * copied and processed in chunks so that the host-to-device transfer overlaps the kernels.  iq_in
 * [n][N_RX][dnrp_rx_stride()] and tb_out [n][N_TB_byte] should be pinned (cudaHostAlloc / cudaHostRegister): both are
 * the target of per-chunk asynchronous copies, pageable memory makes every such copy block the host and removes the
 * overlap (the result is the same).  sync_out / pcc_out / n_iter may be NULL.  harq_slot (here, in dnrp_rx_pdc_batch,
 * dnrp_rx_pdc_batch_ex and dnrp_fec_decode_tb): every slot < max_batch and no slot twice in one batch -- two packets on
 * one slot would combine into the same soft buffer concurrently; refused with DNRP_ERR_INVALID.  On any error the call
 * drains its streams and leaves the context usable. */
int dnrp_rx_batch(dnrp_ctx_t* ctx, uint32_t n, const dnrp_cf32_t* iq_in, uint32_t search_length,
                  const uint32_t* harq_slot, const uint8_t* rv, const uint8_t* harq_reset,
                  dnrp_sync_result_t* sync_out, dnrp_pcc_report_t* pcc_out, uint8_t* tb_out,
                  dnrp_pdc_report_t* pdc_out, uint8_t* n_iter);

/* dnrp_rx_batch for IQ in the radio's wire format: iq_in [n][N_RX][dnrp_rx_stride()][2] int16 (I, Q), each sample
 * becomes (I * scale, Q * scale) on the device (scale = 1/32768 reproduces UHD's sc16 -> fc32 conversion, which the
 * reference leaves to UHD on the host CPU: lib/src/radio/hw_usrp.cpp:436-438, cpu_format "fc32", otw_format "sc16").
 * Halves the host-to-device bytes of dnrp_rx_batch; everything after the conversion is the same code path.
 * dnrp_rx_stride() must be a multiple of 4. */
int dnrp_rx_batch_sc16(dnrp_ctx_t* ctx, uint32_t n, const int16_t* iq_in, float scale, uint32_t search_length,
                       const uint32_t* harq_slot, const uint8_t* rv, const uint8_t* harq_reset,
                       dnrp_sync_result_t* sync_out, dnrp_pcc_report_t* pcc_out, uint8_t* tb_out,
                       dnrp_pdc_report_t* pdc_out, uint8_t* n_iter);

/* The same for UHD's 8-bit wire format (stream args otw_format "sc8"): iq_in [n][N_RX][dnrp_rx_stride()][2] int8.  NOT what
 * the reference's radio layer asks its USRP for (hw_usrp.cpp:424,438 set "sc16"): an option for a host link that is the
 * bottleneck at the price of the 8-bit quantisation (about 38 dB below the signal at 12 dB back-off).
 * scale = 1/128 corresponds to UHD's conversion.  dnrp_rx_stride() must be a multiple of 8. */
int dnrp_rx_batch_sc8(dnrp_ctx_t* ctx, uint32_t n, const int8_t* iq_in, float scale, uint32_t search_length,
                      const uint32_t* harq_slot, const uint8_t* rv, const uint8_t* harq_reset,
                      dnrp_sync_result_t* sync_out, dnrp_pcc_report_t* pcc_out, uint8_t* tb_out,
                      dnrp_pdc_report_t* pdc_out, uint8_t* n_iter);

/* test / scope taps (the reference exposes the same data through its JSON export and TCP scope,
 * rx_synced.cpp:363-443): copies an intermediate device buffer of the last call to the host.
 * Returns DNRP_ERR_INVALID if dst_bytes is smaller than the buffer for n packets. */
typedef enum {
    DNRP_TAP_Y = 1,        /* cf32 [n][N_RX][N_DF][N_b_OCC+1] OFDM symbols 1..N_DF after FFT/scale/STO */
    DNRP_TAP_DRS_ZF = 2,   /* cf32 [n][N_RX][N_eff_TX][n_drs][N_b_OCC/4] */
    DNRP_TAP_STF_ZF = 3,   /* cf32 [n][N_RX][N_b_OCC/4] */
    DNRP_TAP_LLR_PCC = 4,  /* int16 [n][196] */
    DNRP_TAP_LLR_PDC = 5,  /* int16 [n][G] */
    DNRP_TAP_WIENER_CLASS = 6, /* uint8 [n][dnrp_tap_bytes/n] class after each DRS symbol */
    DNRP_TAP_TB_OK = 7,        /* uint8 [n] transport block CRC flags of the last PDC decode */
    DNRP_TAP_CHESTIM = 8       /* cf32 [n][N_RX][N_eff_TX][N_b_OCC+1] interpolated channel after the last data field symbol:
                                * what rx_synced_t::get_json exports as ch_<rx>_<ts> (rx_synced.cpp:444-480) */
} dnrp_tap_t;
size_t dnrp_tap_bytes(const dnrp_ctx_t* ctx, int tap, uint32_t n);
int dnrp_tap_read(dnrp_ctx_t* ctx, int tap, uint32_t n, void* dst, size_t dst_bytes);

/* device-resident loop back: copies the n packets produced by the last dnrp_tx_batch into the
 * receiver's IQ windows (TX antenna t -> RX antenna t % N_TX, offset `t0` samples), the "hardware
 * simulator with an ideal channel" of the reference's loopback firmware. */
int dnrp_loopback_tx_to_rx(dnrp_ctx_t* ctx, uint32_t n, uint32_t t0);

/* ---- simulation channel ------------------------------------------------------------------------
 * One packet's wireless channel + receiver noise, the per-packet equivalent of
 * vspace_t::wchannel_get_samples (lib/src/simulation/vspace.cpp:437-512) for one TX and one RX node:
 * every TX antenna is superimposed on every RX antenna (mode 0 "awgn": unit links,
 * channel_awgn.cpp:39-64; mode 1 "flat": one CN(0,1) coefficient per link, redrawn per packet,
 * channel_flat.cpp:47-88), delayed by t0 samples into the RX window, then white noise of absolute level
 * n0_dBFS = -10 log10(net_bandwidth_norm) - snr_dB is added over the whole window (noise.cpp:31-41).
 * Noise and coefficients come from Philox4x32-10 keyed by `seed` (the reference uses srsRAN's
 * generator: realisations differ, statistics do not). */
typedef struct {
    float snr_dB;
    float net_bandwidth_norm;   /* 1 / oversampling in the reference's simulator */
    uint32_t t0;
    uint32_t mode;              /* 0 awgn, 1 flat, 2 doubly selective (tapped delay line, below) */
    uint64_t seed;
    /* mode 2 only: channel_doubly_t "doubly_<pdp_idx>_<tau_rms_ns>_<fD_Hz>" (channel_doubly.cpp:95-135, link.cpp):
     * generic power delay profile 0..2 stretched to tau_rms_ns (0..2000), 40 sinusoids per tap with maximum Doppler
     * fD_Hz (0..2000), evaluated at absolute sample time time_64 + n of the hardware rate */
    uint32_t pdp_idx;
    float tau_rms_ns;
    float fD_Hz;
    int64_t time_64;
    /* large-scale fading and the simulated radio's converters (all 0 = not used, the behaviour before these fields):
     * large_scale = magnitude gain of every link, channel_t::get_large_scale_via_pathloss (channel.cpp:53-73; helper
     * below); clip_limit = hw_simulator_t::symmetric_clipping_amplitude (1.0 in the reference, hw_simulator.hpp:116):
     * I and Q of the TX samples are clipped and rounded to 2 clip_limit / 2^dac_bits before the channel
     * (hw_simulator.cpp:560-566), those of the RX samples -- after the noise -- with adc_bits (hw_simulator.cpp:685-691);
     * bits = 0: clip only */
    float large_scale;
    float clip_limit;
    uint32_t dac_bits, adc_bits;
} dnrp_channel_desc_t;
/* channel_t::get_large_scale_via_pathloss: db2mag(tx_power_ant_0dBFS - fading - rx_power_ant_0dBFS), fading = free space
 * path loss over distance_m at freq_Hz (pathloss.cpp:28-38) or, for same_node != 0, tx_into_rx_leakage_dB.  Host only. */
float dnrp_large_scale_via_pathloss(float tx_power_ant_0dBFS, float rx_power_ant_0dBFS, float distance_m, float freq_Hz,
                                    int same_node, float tx_into_rx_leakage_dB);

/* reads the n packets left on the device by dnrp_tx_batch, writes the receiver's IQ windows (device);
 * iq_rx_out [n][N_RX][dnrp_rx_stride()] receives a host copy unless NULL */
int dnrp_channel_batch(dnrp_ctx_t* ctx, uint32_t n, const dnrp_channel_desc_t* descs, dnrp_cf32_t* iq_rx_out);

/* ---- measurement hook ------------------------------------------------------------------------
 * Re-runs one device stage `reps` times on the data already resident in the context (left there by
 * the previous host-facing call) and returns the average CUDA-event time per repetition in ms,
 * measured on the context's stream.  stage: see dnrp_stage_t. */
typedef enum {
    DNRP_STAGE_FEC_DECODE_TB = 1,
    DNRP_STAGE_FEC_ENCODE = 2,
    DNRP_STAGE_FEC_DECODE_PLCF = 3,
    DNRP_STAGE_RX_CHAIN = 4,   /* rx front end + PCC + PDC on the resident IQ windows and sync reports */
    DNRP_STAGE_TX_CHAIN = 5,   /* FEC encode + OFDM on the resident TB / PLCF / descriptors */
    DNRP_STAGE_CHANNEL = 6,    /* channel on the resident TX IQ and descriptors */
    DNRP_STAGE_SYNC = 7,       /* packet synchronisation on the resident IQ windows, hand-over to the receiver */
    DNRP_STAGE_SYNC_RX_CHAIN = 8 /* 7 followed by 4: raw IQ windows -> transport blocks */
} dnrp_stage_t;
/* ---- frequency-domain stage of the packet search: beta estimate and integer CFO -------------------------------
 * coarse_peak_f_domain_t::process (lib/src/phy/rx/sync/coarse_peak_f_domain.cpp:70-201), which the reference runs between
 * coarse and fine peak and switches at compile time (sync_param.hpp:244 RX_SYNC_PARAM_AUTOCORRELATOR_PEAK_FIND_BETA_
 * THRESHOLD_DB_OR_ASSUME_MAX_OF_RDC -10.0f, :252 ..._FIND_INTEGER_CFO_SEARCH_RANGE_OR_ASSUME_ZERO 4; both commented out,
 * the integer CFO part "#error not implemented").  Here per context at run time, off by default:
 *   beta_threshold_dB < 0: beta search on (the context's beta is the maximum, as the radio device class' in the
 *       reference); >= 0: off, every packet is taken to have the context's beta;
 *   cfo_search_range_subc >= 4: integer CFO search over the multiples of 4 subcarriers inside +- that range (the STF's
 *       fractional estimate covers +-2); the result is added to the sync result's cfo_integer_rad, which the fine peak
 *       search and the receiver already honour.  Own specification, DESIGN.md 5b.
 * Needs N_b_DFT_os <= 4096 and L = M.  dnrp_sync_fdomain_read: results of the last dnrp_sync_batch / dnrp_rx_batch on this
 * context: estimated beta and where the signal sits in subcarriers (cfo_integer_rad = -2 pi x that / N_b_DFT_os). */
int dnrp_sync_set_fdomain(dnrp_ctx_t* ctx, float beta_threshold_dB, uint32_t cfo_search_range_subc);
int dnrp_sync_fdomain_read(dnrp_ctx_t* ctx, uint32_t n, uint32_t* beta_out, int32_t* cfo_subc_out);

/* ---- residual CFO / STO tracking on the DRS ---------------------------------------------------------------------
 * RX_SYNCED_PARAM_CFO_RESIDUAL_BASED_ON_DRS / RX_SYNCED_PARAM_STO_RESIDUAL_BASED_ON_DRS of the reference
 * (lib/include/dectnrp/phy/rx/rx_synced/rx_synced_param.hpp:156-181, both commented out there: "not functional yet";
 * estimator_cfo_t::process_drs lib/src/phy/rx/rx_synced/offsets/estimator_cfo.cpp:31-70, call sites rx_synced.cpp:839-851,
 * the post-correction rx_synced.cpp:1165-1169 is an empty todo).  Here per context at run time, off by default.  Own
 * specification: per packet ONE estimate over all DRS symbols of the phase advance per N_step OFDM symbols
 * -- common to all subcarriers (residual CFO) and per subcarrier (drift of the symbol timing, i.e. a sampling clock
 * offset) --, the DRS pilots and every cell taken back to the phase reference of the first DRS symbol before channel
 * interpolation and detection (PCC and PDC).  Packets with one DRS occurrence per stream are left as they are.
 * dnrp_rx_tracking_read: out [n][2] = (dphi, dpsi) in rad of the last dnrp_rx_pcc_batch / dnrp_rx_batch on this context;
 * residual CFO = dphi / (2 pi N_step (1 + CP/N)) subcarrier spacings. */
int dnrp_rx_set_tracking(dnrp_ctx_t* ctx, int residual_cfo, int residual_sto);
int dnrp_rx_tracking_read(dnrp_ctx_t* ctx, uint32_t n, float* out);

/* the synchronisation stages run with the search_length of the last dnrp_sync_batch / dnrp_rx_batch call on this context */
int dnrp_time_stage(dnrp_ctx_t* ctx, int stage, uint32_t n, uint32_t reps, float* ms_per_rep);

/* Per-kernel device time inside the stages above: while enabled, every launch of the named kernel is
 * bracketed by CUDA events on the context's stream; dnrp_kernel_time_read synchronises and returns the
 * accumulated milliseconds and the launch count since the last reset. */
typedef enum {
    DNRP_KERNEL_TURBO_DECODE_PDC = 0,
    DNRP_KERNEL_RX_FFT = 1,
    DNRP_KERNEL_RX_DEMAP_PDC = 2,
    DNRP_KERNEL_RX_STF = 3,
    DNRP_KERNEL_DERM_PDC = 4,
    DNRP_KERNEL_SYNC = 5,            /* the four synchronisation kernels together */
    DNRP_KERNEL_RESAMPLE = 6,        /* every k_resample launch (TX, sync window, RX packet) */
    DNRP_KERNEL_COUNT = 7
} dnrp_kernel_t;
int dnrp_kernel_time_enable(dnrp_ctx_t* ctx, int enable);
int dnrp_kernel_time_read(dnrp_ctx_t* ctx, int kernel, float* ms_total, uint32_t* launches);
/* number of kernel launches issued on the context's stream since creation (all kernels of this library) */
uint64_t dnrp_launch_count(const dnrp_ctx_t* ctx);

/* ---- multi-GPU: the counter reduction --------------------------------------------------------
 * Packets (and sweep cells) shard across GPUs with no data-path exchange; the only collective is one all-reduce over the
 * experiment's counter vector.  The reference has no multi-device code: its counters are the members of
 * one result_t / worker statistics object in one process (lib/src/upper/loopback/result.cpp:26-103,
 * lib/src/phy/pool/worker_tx_rx.cpp:96-198 stats.rx_pcc_* / rx_pdc_*); these entry points replace "add into the shared object" by ncclAllReduce over
 * NVLink / NVSwitch.  NCCL is bound at run time (dlopen of libnccl.so.2; nothing else in this library needs it) --
 * DNRP_ERR_STATE if it cannot be loaded.
 *
 * One communicator per context.  Multi-process (one rank per GPU, e.g. under torchrun / MPI): rank 0 calls
 * dnrp_comm_unique_id and hands the 128 bytes to the others by any means, then every rank calls dnrp_comm_init_rank (a
 * collective).  Single process, several GPUs (one batch_phy_t per GPU): dnrp_comm_init_all over the
 * contexts, and the reduction through dnrp_counters_allreduce_all (one NCCL group).
 * dnrp_counters_allreduce: counters[n] int64 (host memory), reduced in place over all ranks; op 0 = sum, 1 = min,
 * 2 = max; n <= DNRP_COMM_MAX_COUNTERS; a collective: every rank calls it with the same n and op.  The reduction runs
 * on the context's stream behind whatever was queued there. */
#define DNRP_COMM_ID_BYTES 128
#define DNRP_COMM_MAX_COUNTERS 65536
int dnrp_comm_unique_id(uint8_t id[DNRP_COMM_ID_BYTES]);
int dnrp_comm_init_rank(dnrp_ctx_t* ctx, int n_ranks, int rank, const uint8_t id[DNRP_COMM_ID_BYTES]);
int dnrp_comm_init_all(dnrp_ctx_t* const* ctxs, int n_ctx);
int dnrp_counters_allreduce(dnrp_ctx_t* ctx, int64_t* counters, uint32_t n, int op);
/* counters [n_ctx][n]: row i belongs to ctxs[i]; every row holds the reduced vector afterwards */
int dnrp_counters_allreduce_all(dnrp_ctx_t* const* ctxs, int n_ctx, int64_t* counters, uint32_t n, int op);
int dnrp_comm_destroy(dnrp_ctx_t* ctx);

#ifdef __cplusplus
}
#endif
#endif
