// api_noalloc.cu -- dnrp_fec_decode_plcf_noalloc: the blind PLCF decode (k_derm + k_turbo_decode<grouped>) for callers
// that must not allocate (CUDA graph capture, tight serving loops).  Same kernels and tables as dnrp_fec_decode_plcf;
// the one difference is where the decoder's scratch lives: the soft buffers the de-rate-matcher writes and the turbo
// decoder reads are NOT context workspaces here but the file-scope __device__ array below, resident on every GPU that
// loads this library until the process exits.  No counterpart in the
// reference.
#include <mutex>
#include "ctx.hpp"

using namespace dnrp;

#define DNRP_REQUIRE(cond, msg)             \
    do {                                    \
        if (!(cond)) {                      \
            dnrp::set_last_error(msg);      \
            return DNRP_ERR_INVALID;        \
        }                                   \
    } while (0)

// soft buffers [packet][2 PLCF types][320] int16 (plan_fec.cu: rx_pcc[ty].soft_off = ty * 320): 1280 B per packet
static constexpr uint32_t kNoallocSoftPerPacket = 640;
static __device__ __align__(256) int16_t g_noalloc_soft[(size_t)DNRP_NOALLOC_MAX_PACKETS * kNoallocSoftPerPacket];

// In host code `g_noalloc_soft` is only a handle of the symbol, not a device address: the address the kernels get is
// looked up once per device, with that device current.
static constexpr int kMaxDevices = 64;
static int16_t* g_noalloc_soft_addr[kMaxDevices] = {};
static std::mutex g_noalloc_mutex;

static int noalloc_scratch(int device, int16_t** out) {
    DNRP_REQUIRE(device >= 0 && device < kMaxDevices, "device index out of range");
    std::lock_guard<std::mutex> lk(g_noalloc_mutex);
    if (!g_noalloc_soft_addr[device]) {
        void* p = nullptr;
        DNRP_CUDA(cudaGetSymbolAddress(&p, g_noalloc_soft));   // the current device's instance of the array
        g_noalloc_soft_addr[device] = static_cast<int16_t*>(p);
    }
    *out = g_noalloc_soft_addr[device];
    return DNRP_OK;
}

extern "C" int dnrp_fec_decode_plcf_noalloc(dnrp_ctx_t* c, uint32_t n, const int16_t* llr196_dev, uint8_t* ok_dev,
                                            uint8_t* plcf_dev, uint32_t* mask_dev, uint8_t* n_iter_dev, void* stream) {
    DNRP_REQUIRE(c && llr196_dev && ok_dev && plcf_dev && mask_dev && n_iter_dev, "null argument");
    DNRP_REQUIRE(n > 0, "empty batch");
    if (n > DNRP_NOALLOC_MAX_PACKETS) {
        set_last_error("dnrp_fec_decode_plcf_noalloc: n = " + std::to_string(n) + " exceeds DNRP_NOALLOC_MAX_PACKETS = " +
                       std::to_string(DNRP_NOALLOC_MAX_PACKETS) + " (size of the static scratch array); split the batch");
        return DNRP_ERR_INVALID;
    }
    DNRP_CUDA(cudaSetDevice(c->device));
    int16_t* soft = nullptr;
    int rc = noalloc_scratch(c->device, &soft);
    if (rc) return rc;
    cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : c->stream;

    tdec_params_t p{};
    p.cbs = c->d_rx_pcc.p; p.C = 2;
    p.llr = llr196_dev; p.llr_stride = 196;
    p.gold = c->d_gold_pcc.p; p.gold_bit_off = 0;
    // Every call shares the array and finds the previous call's values in it.  Nothing has to start at zero: with
    // reset = null k_derm treats every packet as a first transmission and overwrites all 3 (K + 4) soft values of both
    // code blocks, which is all the decoder reads; the CRC flags (ok_dev) are written before they are read.
    p.soft = soft; p.soft_stride = kNoallocSoftPerPacket;
    p.cb_crc = ok_dev;
    p.saved = nullptr; p.saved_stride = 0;
    p.slot = nullptr; p.rv = nullptr; p.reset = nullptr;
    p.tb_out = plcf_dev; p.tb_stride = 32;
    p.n_iter = n_iter_dev;
    p.rm_arena = c->d_rm_arena.p; p.xp_arena = c->d_xp_arena.p;
    p.min_it = 1; p.max_it = 5;  // pcc_enc.cpp:39
    p.crc_mode = 1; p.plcf_mask = mask_dev;
    DNRP_CUDA(launch_derm(p, n, 304, st));
    DNRP_CUDA(launch_turbo_decode(p, n, tdec_geometry(c->fec.rx_pcc, 2), st));
    return DNRP_OK;
}
