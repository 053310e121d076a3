// Arguments of the tcgen05 training-MLP kernels (hn_train_tc.cu), filled by hn_train.cu.
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdint>

namespace hn {
namespace tc {

constexpr int MAX_HIDDEN = 4; // per MLP

struct TrainMlpArgs {
	const __half* w_density; const __half* w_rgb;   // fp16 training weights, row-major [out][in] per layer
	int density_hidden, rgb_hidden;
	uint32_t n;                                      // samples
	// forward activations / their gradients
	__half* enc;                                     // [n][32]
	__half* hid_d[MAX_HIDDEN];                       // density hidden [n][64]
	__half* d_out;                                   // [n][16]
	__half* rgb_in;                                  // [n][32] ([16 density out | 16 SH]; the SH half is written by train_encode_kernel)
	__half* hid_r[MAX_HIDDEN];                       // colour hidden [n][64]
	__half* out;                                     // [n][16]
	const __half* g_out;                             // [n][16] dL/d(out), from the loss kernel
	const float* d_sigma;                            // [n]     dL/d(sigma raw)
	__half* g_hid_r[MAX_HIDDEN];                     // [n][64]
	__half* g_dout;                                  // [n][16]
	__half* g_hid_d[MAX_HIDDEN];                     // [n][64]
	__half* g_enc;                                   // [n][32]
};


constexpr int DW_MAX_LAYERS = 8;
struct DwLayer {
	const __half* a;   // [n][64]
	const __half* b;   // [n][nb]
	float* gw;         // [out][in] fp32 gradient of this layer
	int nb;            // 16, 32 or 64
	int transposed;    // 1: D is [in][out]
	int col;           // first accumulator column
};
struct DwArgs {
	DwLayer layer[DW_MAX_LAYERS];
	int n_layers;
	uint32_t n;
};


} // namespace tc

void launch_train_mlp_forward(const tc::TrainMlpArgs& a, cudaStream_t s);
void launch_train_mlp_backward(const tc::TrainMlpArgs& a, cudaStream_t s);
void launch_train_dw(const tc::DwArgs& a, cudaStream_t s);

} // namespace hn
