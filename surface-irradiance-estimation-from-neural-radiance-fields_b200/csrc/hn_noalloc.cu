// hn_op_encode_noalloc: the hash-grid encoding operator (a-10, hn_op_encode) for callers that must not allocate device memory
// (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc). No counterpart in the
// reference.
//
// hn_op_encode stages the positions and the fp16 features through growable device buffers of the context (cudaMalloc when a call
// is larger than any before). Here the staging lives in ONE file-scope `static __device__` array of fixed size:
//   * every GPU that loads libhn_b200.so holds HN_NOALLOC_SCRATCH_BYTES (4.75 MiB) of it until the process exits;
//   * at most HN_NOALLOC_MAX_POSITIONS (65 536) positions per call, larger calls are rejected with an error;
//   * every call and every context on a device share the array and it keeps the last call's contents. Nothing in it has to start
//     at zero: encode_kernel writes all 32 features of the n positions of the call before they are copied out, and nothing beyond
//     them is read. Calls that overlap on the GPU (two streams, two threads, two contexts on one device) RACE on it - serialise them;
//   * the entry point calls no cudaMalloc / cudaFree / allocating library routine: two copies and one kernel launch on ctx's stream.
//
// In host code the array's name is a host-side stand-in, not a device address. Its device address is looked up once per device,
// with that device current, by cudaGetSymbolAddress in this translation unit, and passed to the kernel as a plain pointer.
#include "hn_context.h"
#include <mutex>

namespace {

constexpr uint32_t MAX_POSITIONS = 65536;                                  // HN_NOALLOC_MAX_POSITIONS in include/hybrid_nerf.h
constexpr size_t POS_BYTES = (size_t)MAX_POSITIONS * 3 * sizeof(float);    // [n][3] fp32 positions   (768 KiB, a multiple of 256)
constexpr size_t ENC_BYTES = (size_t)MAX_POSITIONS * 32 * sizeof(__half);  // [n][32] fp16 features   (4 MiB)
constexpr size_t SCRATCH_BYTES = POS_BYTES + ENC_BYTES;
static_assert(POS_BYTES % 256 == 0, "the feature block must stay 256-byte aligned");

static __device__ __align__(256) unsigned char g_noalloc_scratch[SCRATCH_BYTES];

constexpr int MAX_DEVICES = 64;
std::mutex g_mutex;
unsigned char* g_scratch_ptr[MAX_DEVICES] = {}; // device address of g_noalloc_scratch on each device, nullptr = not looked up yet

// device address of the scratch array on `device` (which must be the current device)
unsigned char* scratch_on_device(int device) {
	if (device < 0 || device >= MAX_DEVICES) throw std::runtime_error("hn_op_encode_noalloc: device ordinal out of range");
	std::lock_guard<std::mutex> lock(g_mutex);
	if (!g_scratch_ptr[device]) {
		void* p = nullptr;
		const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
		if (e != cudaSuccess || !p) throw hn::CudaError(std::string("cudaGetSymbolAddress(g_noalloc_scratch) failed: ") + cudaGetErrorString(e));
		g_scratch_ptr[device] = (unsigned char*)p;
	}
	return g_scratch_ptr[device];
}

} // namespace

extern "C" int hn_op_encode_noalloc(hn_context* ctx, const float* positions, uint32_t n, uint16_t* out) {
	if (!ctx) return 1;
	try {
		if (!ctx->has_params) throw std::runtime_error("network parameters not set");
		if (n > MAX_POSITIONS) throw std::runtime_error("hn_op_encode_noalloc: n = " + std::to_string(n) + " positions exceed the static scratch (HN_NOALLOC_MAX_POSITIONS = " + std::to_string(MAX_POSITIONS) + "); split the call or use hn_op_encode");
		if (!n) return 0;
		if (!positions || !out) throw std::runtime_error("hn_op_encode_noalloc: null input or output pointer");
		HN_CUDA(cudaSetDevice(ctx->device));
		unsigned char* base = scratch_on_device(ctx->device);
		float* d_p = reinterpret_cast<float*>(base);
		__half* d_o = reinterpret_cast<__half*>(base + POS_BYTES);
		HN_CUDA(cudaMemcpyAsync(d_p, positions, (size_t)n * 3 * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
		hn::launch_encode(ctx->levels, ctx->net.F, ctx->params.p + ctx->net.grid_offset, d_p, n, d_o, ctx->stream);
		HN_CUDA(cudaGetLastError());
		HN_CUDA(cudaMemcpyAsync(out, d_o, (size_t)n * 32 * sizeof(__half), cudaMemcpyDeviceToHost, ctx->stream));
		HN_CUDA(cudaStreamSynchronize(ctx->stream));
		return 0;
	} catch (const std::exception& e) { ctx->error = e.what(); return 1; } catch (...) { ctx->error = "unknown error"; return 1; }
}
