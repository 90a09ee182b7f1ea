// arachne_ksw_extend_static(): the ksw_extend2 batch (k_ext_dp, ext_stage.cuh) for callers that cannot call cudaMalloc -- code under
// CUDA graph capture, tight serving loops, harnesses that forbid allocation.  No counterpart in the reference.
// Included at the end of pipeline.cu (same translation unit as the kernels and as the array below).
//
// Everything k_ext_dp needs besides the caller's inputs and outputs -- the length-sorted problem queue, the per-problem results, the
// queue head / counters -- lives in ONE file-scope device array of fixed size.  Every GPU that loads the library holds the array until
// the process exits.  All calls share it and it keeps the last call's contents: the control block is cleared at the start of every
// call, the queue / result slots a call reads are all written by that call first.  Two calls that overlap on the GPU (different streams,
// or two host threads) race on it; calls on one stream are ordered and safe.
//
// In HOST code the array's name is a host-side stand-in, not a device address; the address the kernels get comes from
// cudaGetSymbolAddress(), asked once per device with that device current.

#define ARACHNE_XS_MAX 65536  // == ARACHNE_KSW_STATIC_MAX_PROBLEMS (include/arachne_cuda.h)
#include "xs_kernels.cuh"  // XsCtl, XsIn, k_xs_count / k_xs_base / k_xs_queue / k_xs_finish

namespace arachne {

constexpr size_t XS_CTL_BYTES = 8192;
static_assert(sizeof(XsCtl) <= XS_CTL_BYTES, "control block");
constexpr size_t XS_QUEUE_OFF = XS_CTL_BYTES;
constexpr size_t XS_RES_OFF = XS_QUEUE_OFF + (size_t)ARACHNE_XS_MAX * sizeof(ExtProb);
constexpr size_t XS_BYTES = XS_RES_OFF + (size_t)ARACHNE_XS_MAX * sizeof(ExtRes);  // 8 KB + 3 MB + 1.5 MB

}  // namespace arachne

static __device__ __align__(256) unsigned char g_xs_scratch[arachne::XS_BYTES];

namespace arachne {

struct XsDevice { bool ready = false; unsigned char* scratch = nullptr; int sms = 0; int per_sm[ARACHNE_MAX_READ_LEN + 1]; };
static XsDevice g_xs_dev[64];
static std::mutex g_xs_mu;

}  // namespace arachne

extern "C" int arachne_ksw_extend_static(int device, const arachne_mem_opt_t* opt_, int64_t n, const uint8_t* d_q, const int64_t* d_q_off,
                                         const uint8_t* d_t, const int64_t* d_t_off, const int32_t* d_h0, const int32_t* d_w, int max_qlen,
                                         int end_bonus, int zdrop, int32_t* d_out, int32_t* d_status, uint64_t* d_cells, void* cuda_stream) {
    using namespace arachne;
    if (!d_q || !d_q_off || !d_t || !d_t_off || !d_h0 || !d_w || !d_out || !d_status || n < 0) { set_error("arachne_ksw_extend_static: bad argument"); return ARACHNE_ERR_ARG; }
    if (n > ARACHNE_XS_MAX) {
        set_error("arachne_ksw_extend_static: %lld problems; the static scratch holds at most %d per call (split the batch, or use arachne_ksw_extend_batch)",
                  (long long)n, ARACHNE_XS_MAX);
        return ARACHNE_ERR_ARG;
    }
    if (max_qlen < 1 || max_qlen > ARACHNE_MAX_READ_LEN) { set_error("arachne_ksw_extend_static: max_qlen must be 1..%d", ARACHNE_MAX_READ_LEN); return ARACHNE_ERR_ARG; }
    int rc = validate_opts(opt_);
    if (rc) return rc;
    if (device < 0 || device >= 64 || arachne_device_count() <= device) { set_error("no CUDA device %d (no CPU path)", device); return ARACHNE_ERR_CUDA; }
    AR_CUDA(cudaSetDevice(device));
    auto dp_kernel = k_ext_dp<false, true>;  // the 13-bit column codec: any h0 + qlen * a <= 8191, no look at the inputs needed on the host
    const int ncols = max_qlen + 2;
    const size_t smem = (size_t)ncols * XDP_BLOCK * sizeof(uint32_t);
    XsDevice& D = g_xs_dev[device];
    {
        // set-up, once per device, with that device current: the device address of the scratch array and the launch shape
        std::lock_guard<std::mutex> lk(g_xs_mu);
        if (!D.ready) {
            void* p = nullptr;
            AR_CUDA(cudaGetSymbolAddress(&p, g_xs_scratch));
            D.scratch = static_cast<unsigned char*>(p);
            D.sms = sm_count(device);
            for (int l = 0; l <= ARACHNE_MAX_READ_LEN; ++l) D.per_sm[l] = 0;
            D.ready = true;
        }
        // every call: arachne_ksw_extend_batch sets the same attribute of the same kernel to ITS batch's size (not an allocation)
        AR_CUDA(cudaFuncSetAttribute(dp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)((size_t)(ARACHNE_MAX_READ_LEN + 2) * XDP_BLOCK * sizeof(uint32_t))));
        if (D.per_sm[max_qlen] == 0) {
            int per_sm = 1;
            AR_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, dp_kernel, XDP_BLOCK, smem));
            D.per_sm[max_qlen] = per_sm < 1 ? 1 : per_sm;
        }
    }
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    XsCtl* ctl = reinterpret_cast<XsCtl*>(D.scratch);
    ExtProb* queue = reinterpret_cast<ExtProb*>(D.scratch + XS_QUEUE_OFF);
    ExtRes* res = reinterpret_cast<ExtRes*>(D.scratch + XS_RES_OFF);
    MemOpt opt; memcpy(&opt, opt_, sizeof(opt));
    XsIn I; I.n = (long)n; I.q_off = d_q_off; I.t_off = d_t_off; I.h0 = d_h0; I.w = d_w; I.end_bonus = end_bonus; I.max_qlen = max_qlen; I.a = opt.a;
    const int grid = (int)std::min<int64_t>((n + 255) / 256 + 1, (int64_t)D.sms * 8);
    AR_CUDA(cudaMemsetAsync(ctl, 0, sizeof(XsCtl), st));  // the array keeps the last call's contents: bins, cursors, head, counters start at zero
    k_xs_count<<<grid, 256, 0, st>>>(I, ctl, d_out);
    k_xs_base<<<1, 256, 0, st>>>(ctl);
    k_xs_queue<<<grid, 256, 0, st>>>(I, ctl, queue);
    ExtDpArgs DA;
    DA.seq = d_q; DA.tbytes = d_t; DA.queue = queue; DA.res = res; DA.n_items = &ctl->n_items; DA.head = &ctl->head;
    DA.ncols = ncols; DA.zdrop = zdrop; DA.one = 1; DA.cnt = &ctl->cnt;
    RefView no_ref; memset(&no_ref, 0, sizeof(no_ref));
    dp_kernel<<<D.sms * D.per_sm[max_qlen], XDP_BLOCK, smem, st>>>(opt, no_ref, DA);
    k_xs_finish<<<grid, 256, 0, st>>>(I, ctl, res, d_out, d_status, reinterpret_cast<unsigned long long*>(d_cells));
    AR_CUDA(cudaGetLastError());
    return ARACHNE_OK;
}
