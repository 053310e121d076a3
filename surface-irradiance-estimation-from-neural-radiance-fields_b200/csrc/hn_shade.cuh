// Activations and colour-space helpers shared by the compositing code (hn_composite.cu and the fused epilogue of
// hn_network_tc.cu): nerf_device.cuh:203-263 (network_to_rgb/density), common_device.cuh srgb<->linear.
#pragma once
#include "hn_internal.h"

namespace hn {

#define HN_MIN_STEP (1.73205080757f / 1024.0f)

__device__ __forceinline__ float act_rgb(float v, int act) {
	switch (act) {
		case 1: return fmaxf(v, 0.0f);
		case 2: return 1.0f / (1.0f + __expf(-v));
		case 3: return __expf(fminf(fmaxf(v, -10.0f), 10.0f));
		default: return v;
	}
}
__device__ __forceinline__ float act_density(float v, int act) {
	switch (act) {
		case 1: return fmaxf(v, 0.0f);
		case 2: return 1.0f / (1.0f + __expf(-v));
		case 3: return __expf(v);
		default: return v;
	}
}
__device__ __forceinline__ float srgb_to_linear(float s) { return s <= 0.04045f ? s / 12.92f : powf((s + 0.055f) / 1.055f, 2.4f); }
__device__ __forceinline__ float linear_to_srgb(float l) { return l < 0.0031308f ? 12.92f * l : 1.055f * powf(l, 0.41666f) - 0.055f; }

} // namespace hn
