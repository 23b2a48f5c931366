// fec_rx.cuh -- launch interface of the RX FEC kernels (fec_rx.cu)
#pragma once
#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>
#include "turbo_dec.cuh"

namespace dnrp {

struct tdec_params_t {
    const cb_desc_t* cbs;      // [C] code block descriptors of the plan (device)
    uint32_t C;
    const int16_t* llr;        // [packet][llr_stride] demapper output (still scrambled)
    size_t llr_stride;
    const uint8_t* gold;       // table of packed scrambling sequences (MSB first), gold_stride bytes each
    uint32_t gold_bit_off;
    const uint32_t* gold_sel;  // [packet] sequence of the table (network id / PLCF type of the packet; null: sequence 0)
    uint32_t gold_stride;
    int16_t* soft;             // [slot][soft_stride] HARQ soft buffers, d-stream layout [3][K+4] per CB
    size_t soft_stride;
    uint8_t* cb_crc;           // [slot][C]
    uint8_t* saved;            // [slot][saved_stride] decoded bytes of CBs whose CRC passed (may be null)
    size_t saved_stride;
    const uint32_t* slot;      // [packet] HARQ slot (null: slot = packet)
    const uint8_t* rv;         // [packet] redundancy version (null: 0)
    const uint8_t* reset;      // [packet] 1 = first transmission, ignore old soft values (null: 1)
    uint8_t* tb_out;           // [packet][tb_stride]
    size_t tb_stride;
    uint8_t* n_iter;           // [packet][C]
    const uint16_t* rm_arena;
    const uint32_t* xp_arena;
    uint32_t min_it, max_it;
    uint32_t crc_mode;         // 0: remainder 0 (PDC), 1: remainder in {0,5555,AAAA,FFFF} (PLCF)
    uint32_t* plcf_mask;       // [packet][C] matched mask or 0xffffffff (crc_mode 1)
};

struct tbfin_params_t {
    uint32_t C, tbs_bits, chunk_bytes;
    const uint32_t* slot;
    uint8_t* cb_crc;
    const uint8_t* tb_out;
    size_t tb_stride;
    const uint32_t* xp_tb;     // [nchunk] x^(8*bytes after chunk i) mod CRC24A
    uint8_t* tb_ok;            // [packet]
};

size_t tdec_smem_bytes(uint32_t K, uint32_t L, bool ck_alias = false);
// descramble + de-rate-match + HARQ combine into the soft buffers; must precede launch_turbo_decode.
// max_n_e = largest LLR count of a code block in p.cbs (sizes the shared staging buffer)
cudaError_t launch_derm(const tdec_params_t& p, uint32_t n_packets, uint32_t max_n_e, cudaStream_t st);
// thread/shared-memory geometry of a decoder launch over the n descriptors cbs (host copies)
struct tdec_geom_t {
    uint32_t threads, tg;  // block size, threads per (code block, packet) item
    size_t smem_item;      // dynamic shared bytes per item
    bool ck_alias;         // k_turbo_decode<.., CKALIAS = true>: checkpoints 0 and 1 in the AE / BS rows (four blocks per SM)
};
tdec_geom_t tdec_geometry(const cb_desc_t* cbs, uint32_t n);
cudaError_t launch_turbo_decode(const tdec_params_t& p, uint32_t n_packets, const tdec_geom_t& g, cudaStream_t st);
cudaError_t launch_tb_finalize(const tbfin_params_t& p, uint32_t n_packets, cudaStream_t st);

}  // namespace dnrp
