// fec_tx.cuh -- launch interface of the TX FEC kernel (fec_tx.cu)
#pragma once
#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>

namespace dnrp {

struct txcb_desc_t {
    uint32_t K, f1, f2;
    uint32_t rp, n_data;       // first payload bit inside the input message, payload bits of this CB
    uint32_t tb_crc_bits, tb_crc_poly, tb_crc_len;  // whole-message CRC appended to this CB (0 = none)
    uint32_t cb_crc;           // 1: append CRC24B over the CB
    uint32_t wp, n_e, n_pad;   // output bit offset, rate matched bits, extra scrambled zero bits
    uint32_t rm_src_off;       // offset of this K's src[] table in the rate-matching arena
    uint32_t rm_start[4];
    uint32_t Nc;               // usable entries of src[] (3K + 12 unless the soft buffer is limited)
    // CRC chunk combination: x^(bits behind thread t's chunk + order) mod g for the TX_FEC_THREADS chunks of the
    // whole-message CRC (xp_tb_off) and of the code block CRC (xp_cb_off), offsets into the x^n arena
    uint32_t xp_tb_off, xp_cb_off;
};
constexpr uint32_t TX_FEC_THREADS = 256;

struct txfec_params_t {
    const txcb_desc_t* cbs;
    uint32_t C;
    const uint8_t* in;         // [packet][in_stride] packed message bytes (TB or PLCF)
    size_t in_stride;
    const uint8_t* rv;         // [packet] or null
    const uint32_t* plcf_mask; // [packet] CRC mask (PLCF only) or null
    const uint8_t* gold;       // table of packed scrambling sequences, gold_stride bytes each
    uint32_t gold_bit_off;
    const uint32_t* gold_sel;  // [packet] sequence of the table this packet is scrambled with (null: sequence 0)
    uint32_t gold_stride;
    const uint8_t* cb_sel;     // [packet] offset added to the code block index (PLCF: 0 = type 1, 1 = type 2; null: 0)
    uint8_t* out;              // [packet][out_stride] packed coded bits, zero-initialised, stride % 4 == 0
    size_t out_stride;
    const uint16_t* rm_arena;
    const uint32_t* xp_arena;
};

cudaError_t launch_tx_fec(const txfec_params_t& p, uint32_t n_packets, uint32_t Kmax, cudaStream_t st);

}  // namespace dnrp
