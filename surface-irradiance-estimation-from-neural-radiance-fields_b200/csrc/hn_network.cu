// (b)+(c) Fused network evaluation for sm_100a: multiresolution hash-grid gather -> density MLP -> SH -> colour MLP in ONE
// kernel; the 64 B/sample encoded features and all hidden activations never touch global memory.
// Samples arrive in ray chunks of CHUNK=32 (one warp = one ray chunk = one direction). Output: 8 B/sample (r,g,b,sigma fp16).
//
// Arithmetic follows the reference (tcnn grid.h:48-212, common_device.h:340-366,651-718,842-854; fully_fused_mlp.cu):
// corner weights rounded to fp16 and accumulated with fp16 FMAs (grid.h:162) -> the encoding is bit-identical to the oracle;
// MLPs take fp16 inputs/weights, accumulate in fp32 on the tensor cores and store fp16 activations between layers.
#include "hn_network.cuh"
#include <mma.h>

namespace hn {

__device__ __forceinline__ void ldmatrix_x4(uint32_t (&r)[4], uint32_t addr) {
	asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];\n" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
	asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
		: "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}


// operator entry point a-10: encoding only
template <int F>
__global__ void __launch_bounds__(128) encode_kernel(GridLevels g, const __half* __restrict__ table, const float* __restrict__ pos, uint32_t n, __half* __restrict__ out) {
	uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	uint32_t enc[16];
	encode_sample<F>(g, table, pos[3 * i], pos[3 * i + 1], pos[3 * i + 2], enc);
	uint4* o = reinterpret_cast<uint4*>(out + (size_t)i * 32);
	#pragma unroll
	for (int k = 0; k < 4; ++k) o[k] = make_uint4(enc[4 * k], enc[4 * k + 1], enc[4 * k + 2], enc[4 * k + 3]);
}


// ---- MLP on legacy-free warp MMA tiles: each warp owns 32 rows (2 m-tiles of 16); activations stay in registers.
// Weights live in shared memory as [out][in] rows padded by 16 B (conflict-free ldmatrix); a row of W is one B column.
constexpr int W32_STRIDE = 80;   // bytes per weight row with 32 inputs
constexpr int W64_STRIDE = 144;  // bytes per weight row with 64 inputs
constexpr int A_STRIDE = 80;     // bytes per staged input row (32 halfs + pad)

// C[2][NT][4] += A[2][KT][4] x W^T, W rows = NT*8 outputs
template <int KT, int NT, int STRIDE>
__device__ __forceinline__ void mma_layer(const uint32_t (&A)[2][KT][4], uint32_t w_smem, float (&C)[2][NT][4], int lane) {
	#pragma unroll
	for (int mt = 0; mt < 2; ++mt)
		#pragma unroll
		for (int nt = 0; nt < NT; ++nt)
			#pragma unroll
			for (int k = 0; k < 4; ++k) C[mt][nt][k] = 0.0f;
	const int mid = lane >> 3, row = lane & 7;
	#pragma unroll
	for (int kt = 0; kt < KT; ++kt) {
		#pragma unroll
		for (int np = 0; np < NT / 2; ++np) {
			uint32_t b[4];
			uint32_t addr = w_smem + (uint32_t)(((2 * np + (mid >> 1)) * 8 + row) * STRIDE + (kt * 16 + (mid & 1) * 8) * 2);
			ldmatrix_x4(b, addr);
			#pragma unroll
			for (int mt = 0; mt < 2; ++mt) {
				mma16816(C[mt][2 * np], A[mt][kt], b[0], b[1]);
				mma16816(C[mt][2 * np + 1], A[mt][kt], b[2], b[3]);
			}
		}
	}
}
// fp32 accumulators of 64 outputs -> fp16 A fragments (4 k-tiles) of the next layer, with ReLU
__device__ __forceinline__ void relu_to_a(const float (&C)[2][8][4], uint32_t (&A)[2][4][4]) {
	#pragma unroll
	for (int mt = 0; mt < 2; ++mt)
		#pragma unroll
		for (int kt = 0; kt < 4; ++kt) {
			A[mt][kt][0] = pack_h2(fmaxf(C[mt][2 * kt][0], 0.f), fmaxf(C[mt][2 * kt][1], 0.f));
			A[mt][kt][1] = pack_h2(fmaxf(C[mt][2 * kt][2], 0.f), fmaxf(C[mt][2 * kt][3], 0.f));
			A[mt][kt][2] = pack_h2(fmaxf(C[mt][2 * kt + 1][0], 0.f), fmaxf(C[mt][2 * kt + 1][1], 0.f));
			A[mt][kt][3] = pack_h2(fmaxf(C[mt][2 * kt + 1][2], 0.f), fmaxf(C[mt][2 * kt + 1][3], 0.f));
		}
}

__host__ __device__ inline uint32_t mlp_smem_bytes(int hidden) { return 64 * W32_STRIDE + (hidden - 1) * 64 * W64_STRIDE + 16 * W64_STRIDE; }

// copy one MLP's weights (global, dense rows) into padded shared memory rows
__device__ void stage_mlp(const __half* __restrict__ w, int hidden, uint8_t* smem) {
	// first layer: 64 rows x 32 halfs (4 x 16 B per row)
	for (int i = threadIdx.x; i < 64 * 4; i += blockDim.x) {
		int r = i >> 2, c = i & 3;
		*reinterpret_cast<uint4*>(smem + r * W32_STRIDE + c * 16) = __ldg(reinterpret_cast<const uint4*>(w + r * 32) + c);
	}
	w += 64 * 32; smem += 64 * W32_STRIDE;
	int rows64 = (hidden - 1) * 64 + 16;
	for (int i = threadIdx.x; i < rows64 * 8; i += blockDim.x) {
		int r = i >> 3, c = i & 7;
		*reinterpret_cast<uint4*>(smem + r * W64_STRIDE + c * 16) = __ldg(reinterpret_cast<const uint4*>(w + r * 64) + c);
	}
}

// hidden stack of one MLP after its 32-wide input layer: returns the 16-wide output accumulators
__device__ __forceinline__ void mlp_forward(const uint32_t (&Ain)[2][2][4], uint32_t w_smem, int hidden, float (&Cout)[2][2][4], int lane) {
	float C[2][8][4];
	uint32_t A[2][4][4];
	mma_layer<2, 8, W32_STRIDE>(Ain, w_smem, C, lane);
	relu_to_a(C, A);
	w_smem += 64 * W32_STRIDE;
	for (int l = 1; l < hidden; ++l) {
		mma_layer<4, 8, W64_STRIDE>(A, w_smem, C, lane);
		relu_to_a(C, A);
		w_smem += 64 * W64_STRIDE;
	}
	mma_layer<4, 2, W64_STRIDE>(A, w_smem, Cout, lane);
}

template <int F>
__global__ void __launch_bounds__(128) eval_kernel(const __grid_constant__ EvalArgs a) {
	extern __shared__ __align__(16) uint8_t smem[];
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	uint8_t* w_density_s = smem;
	uint8_t* w_rgb_s = w_density_s + mlp_smem_bytes(a.density_hidden);
	uint8_t* stage = w_rgb_s + mlp_smem_bytes(a.rgb_hidden) + warp * (32 * A_STRIDE);
	stage_mlp(a.w_density, a.density_hidden, w_density_s);
	stage_mlp(a.w_rgb, a.rgb_hidden, w_rgb_s);
	__syncthreads();

	const uint32_t n_chunks = a.n_chunks_ptr ? min(*a.n_chunks_ptr, a.n_chunks) : a.n_chunks;
	const uint32_t warps_total = gridDim.x * (blockDim.x >> 5);
	const int g = lane >> 2, tg = lane & 3;

	for (uint32_t chunk = blockIdx.x * (blockDim.x >> 5) + warp; chunk < n_chunks; chunk += warps_total) {
		const uint32_t count = a.sample_counts ? a.sample_counts[chunk] : CHUNK;
		if (count == 0) continue;
		// ---- (b) gather: lane = sample row
		{
			uint32_t enc[16];
			#pragma unroll
			for (int k = 0; k < 16; ++k) enc[k] = 0u;
			if ((uint32_t)lane < count) {
				float4 s = __ldg(a.samples + (size_t)chunk * CHUNK + lane);
				// warp_position (nerf_device.cuh:265-270): (p - min) / diag, IEEE divide
				float px = __fdiv_rn(__fsub_rn(s.x, a.train_lo[0]), a.train_diag[0]);
				float py = __fdiv_rn(__fsub_rn(s.y, a.train_lo[1]), a.train_diag[1]);
				float pz = __fdiv_rn(__fsub_rn(s.z, a.train_lo[2]), a.train_diag[2]);
				encode_sample<F>(a.g, a.table, px, py, pz, enc);
			}
			uint4* row = reinterpret_cast<uint4*>(stage + lane * A_STRIDE);
			#pragma unroll
			for (int k = 0; k < 4; ++k) row[k] = make_uint4(enc[4 * k], enc[4 * k + 1], enc[4 * k + 2], enc[4 * k + 3]);
		}
		__syncwarp();
		// ---- (c) density MLP
		uint32_t Ain[2][2][4];
		{
			const int mid = lane >> 3, r = lane & 7;
			#pragma unroll
			for (int mt = 0; mt < 2; ++mt)
				#pragma unroll
				for (int kt = 0; kt < 2; ++kt)
					ldmatrix_x4(Ain[mt][kt], smem_u32(stage + (mt * 16 + (mid & 1) * 8 + r) * A_STRIDE + (kt * 16 + (mid >> 1) * 8) * 2));
		}
		__syncwarp();
		float Cd[2][2][4];
		mlp_forward(Ain, smem_u32(w_density_s), a.density_hidden, Cd, lane);
		// ---- colour MLP input: [16 density outputs | 16 SH coefficients]
		float dx, dy, dz;
		if (a.dirs_override) {
			dx = a.dirs_override[3 * chunk]; dy = a.dirs_override[3 * chunk + 1]; dz = a.dirs_override[3 * chunk + 2];
		} else {
			float4 d = __ldg(a.d_mw + a.alive_list[chunk]);
			dx = d.x; dy = d.y; dz = d.z;
		}
		// warp_direction then kernel_sh's *2-1 (nerf_device.cuh:290-292, spherical_harmonics.h:66-68)
		dx = __fmul_rn(__fadd_rn(dx, 1.0f), 0.5f) * 2.f - 1.f;
		dy = __fmul_rn(__fadd_rn(dy, 1.0f), 0.5f) * 2.f - 1.f;
		dz = __fmul_rn(__fadd_rn(dz, 1.0f), 0.5f) * 2.f - 1.f;
		float sh[16];
		sh4(dx, dy, dz, sh);
		uint32_t p0, p1;
		if (tg == 0) { p0 = pack_h2(sh[0], sh[1]); p1 = pack_h2(sh[8], sh[9]); }
		else if (tg == 1) { p0 = pack_h2(sh[2], sh[3]); p1 = pack_h2(sh[10], sh[11]); }
		else if (tg == 2) { p0 = pack_h2(sh[4], sh[5]); p1 = pack_h2(sh[12], sh[13]); }
		else { p0 = pack_h2(sh[6], sh[7]); p1 = pack_h2(sh[14], sh[15]); }
		__half sigma[2][2];
		#pragma unroll
		for (int mt = 0; mt < 2; ++mt) {
			Ain[mt][0][0] = pack_h2(Cd[mt][0][0], Cd[mt][0][1]);
			Ain[mt][0][1] = pack_h2(Cd[mt][0][2], Cd[mt][0][3]);
			Ain[mt][0][2] = pack_h2(Cd[mt][1][0], Cd[mt][1][1]);
			Ain[mt][0][3] = pack_h2(Cd[mt][1][2], Cd[mt][1][3]);
			Ain[mt][1][0] = p0; Ain[mt][1][1] = p0; Ain[mt][1][2] = p1; Ain[mt][1][3] = p1;
			sigma[mt][0] = __float2half_rn(Cd[mt][0][0]); // density channel 0, rows g / g+8 (valid on tg == 0)
			sigma[mt][1] = __float2half_rn(Cd[mt][0][2]);
		}
		float Cr[2][2][4];
		mlp_forward(Ain, smem_u32(w_rgb_s), a.rgb_hidden, Cr, lane);
		// ---- output: tg 0 holds (r,g), tg 1 holds (b,.) of rows g and g+8
		#pragma unroll
		for (int mt = 0; mt < 2; ++mt) {
			float b_lo = __shfl_down_sync(0xffffffffu, Cr[mt][0][0], 1);
			float b_hi = __shfl_down_sync(0xffffffffu, Cr[mt][0][2], 1);
			if (tg == 0) {
				uint2 o0, o1;
				o0.x = pack_h2(Cr[mt][0][0], Cr[mt][0][1]);
				o1.x = pack_h2(Cr[mt][0][2], Cr[mt][0][3]);
				__half2 t0 = __halves2half2(__float2half_rn(b_lo), sigma[mt][0]);
				__half2 t1 = __halves2half2(__float2half_rn(b_hi), sigma[mt][1]);
				o0.y = *reinterpret_cast<uint32_t*>(&t0);
				o1.y = *reinterpret_cast<uint32_t*>(&t1);
				a.out[(size_t)chunk * CHUNK + mt * 16 + g] = o0;
				a.out[(size_t)chunk * CHUNK + mt * 16 + g + 8] = o1;
			}
		}
	}
}

// ---------------------------------------------------------------- launchers
void launch_encode(const GridLevels& g, uint32_t F, const __half* table, const float* positions, uint32_t n, __half* out, cudaStream_t s) {
	if (!n) return;
	if (F == 4) encode_kernel<4><<<(n + 127) / 128, 128, 0, s>>>(g, table, positions, n, out);
	else encode_kernel<2><<<(n + 127) / 128, 128, 0, s>>>(g, table, positions, n, out);
}

// ---- ERenderMode::Normals (src/testbed_nerf.cu:2106-2107,688-693): NerfNetwork::input_gradient(dim 3) = gradient of the raw
// density output with respect to the warped position, through the density MLP and the hash grid. A visualisation mode, so
// plain CUDA cores and fp32: one thread per sample, density-MLP weights staged in shared memory.
//   forward : enc (the bit-exact fp16 encoding) -> hidden layers (fp32 accumulate, fp16 activations as in inference)
//   backward: one-hot on output 0 -> W_out row 0 masked by ReLU -> ... -> W_first^T -> 32 encoding gradients
//   grid    : d enc[l,f] / d x_d = scale_l * sum over the 4 corner pairs along d of w_other * (v_hi - v_lo)   (grid.h:170-212)
// The sample's (r,g,b) slots of the network output are overwritten with normalize(-dsigma/draw * grad) as fp16.
template <int F>
__global__ void __launch_bounds__(128) normals_kernel(GridLevels g, const __half* __restrict__ table, const __half* __restrict__ w_density, int hidden_layers,
	float3 lo, float3 diag, const float4* __restrict__ samples, const uint32_t* __restrict__ sample_counts, const uint32_t* __restrict__ n_chunks_ptr,
	uint32_t n_chunks, int density_act, uint2* __restrict__ out) {
	extern __shared__ __half w_s[];
	const int n_w = 64 * 32 + (hidden_layers - 1) * 64 * 64 + 16 * 64;
	for (int i = threadIdx.x; i < n_w; i += blockDim.x) w_s[i] = w_density[i];
	__syncthreads();
	const uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
	const uint32_t chunk = idx >> 5, lane = idx & 31u;
	uint32_t nc = n_chunks;
	if (n_chunks_ptr) nc = min(nc, *n_chunks_ptr);
	if (chunk >= nc) return;
	const uint32_t count = sample_counts ? sample_counts[chunk] : CHUNK;
	if (lane >= count) return;
	const float4 s = __ldg(samples + idx);
	const float px = __fdiv_rn(__fsub_rn(s.x, lo.x), diag.x), py = __fdiv_rn(__fsub_rn(s.y, lo.y), diag.y), pz = __fdiv_rn(__fsub_rn(s.z, lo.z), diag.z);
	uint32_t encu[16];
	encode_sample<F>(g, table, px, py, pz, encu);
	float enc[32];
	#pragma unroll
	for (int k = 0; k < 16; ++k) { const __half2 h = *reinterpret_cast<const __half2*>(&encu[k]); enc[2 * k] = __low2float(h); enc[2 * k + 1] = __high2float(h); }
	// forward through the hidden layers (at most 2, checked by the launcher)
	float h0[64], h1[64];
	const __half* w = w_s;
	for (int o = 0; o < 64; ++o) {
		float acc = 0.0f;
		#pragma unroll
		for (int i = 0; i < 32; ++i) acc += __half2float(w[o * 32 + i]) * enc[i];
		h0[o] = __half2float(__float2half_rn(fmaxf(acc, 0.0f)));
	}
	w += 64 * 32;
	if (hidden_layers == 2) {
		for (int o = 0; o < 64; ++o) {
			float acc = 0.0f;
			for (int i = 0; i < 64; ++i) acc += __half2float(w[o * 64 + i]) * h0[i];
			h1[o] = __half2float(__float2half_rn(fmaxf(acc, 0.0f)));
		}
		w += 64 * 64;
	}
	// raw density (output row 0) and the backward pass of its one-hot gradient
	float raw = 0.0f;
	float gl[64]; // gradient at the last hidden layer
	{
		const float* hl = hidden_layers == 2 ? h1 : h0;
		for (int i = 0; i < 64; ++i) { const float wv = __half2float(w[i]); raw += wv * hl[i]; gl[i] = hl[i] > 0.0f ? wv : 0.0f; }
	}
	raw = __half2float(__float2half_rn(raw));
	if (hidden_layers == 2) {
		const __half* w2 = w_s + 64 * 32;
		float gp[64];
		for (int i = 0; i < 64; ++i) {
			float acc = 0.0f;
			for (int o = 0; o < 64; ++o) acc += __half2float(w2[o * 64 + i]) * gl[o];
			gp[i] = h0[i] > 0.0f ? acc : 0.0f;
		}
		for (int i = 0; i < 64; ++i) gl[i] = gp[i];
	}
	float genc[32];
	#pragma unroll
	for (int i = 0; i < 32; ++i) genc[i] = 0.0f;
	for (int o = 0; o < 64; ++o) {
		const float go = gl[o];
		#pragma unroll
		for (int i = 0; i < 32; ++i) genc[i] += __half2float(w_s[o * 32 + i]) * go;
	}
	// hash grid: gradient of every feature with respect to the position
	float gx = 0.0f, gy = 0.0f, gz = 0.0f;
	constexpr int L = 32 / F;
	for (int l = 0; l < L; ++l) {
		const float scale = g.scale[l];
		const uint32_t res = g.res[l], size = g.size[l], mask = g.mask[l];
		const __half* __restrict__ lvl = table + (size_t)g.offset[l] * F;
		float fx = fmaf(scale, px, 0.5f), fy = fmaf(scale, py, 0.5f), fz = fmaf(scale, pz, 0.5f);
		const float flx = floorf(fx), fly = floorf(fy), flz = floorf(fz);
		const uint32_t x0 = (uint32_t)(int)flx, y0 = (uint32_t)(int)fly, z0 = (uint32_t)(int)flz;
		fx -= flx; fy -= fly; fz -= flz;
		const bool hashed = g.hashed[l] != 0;
		float v[8][F];
		#pragma unroll
		for (int c = 0; c < 8; ++c) {
			const uint32_t x = x0 + (c & 1), y = y0 + ((c >> 1) & 1), z = z0 + ((c >> 2) & 1);
			uint32_t i = hashed ? (x ^ (y * 2654435761u) ^ (z * 805459861u)) : (x + y * res + z * res * res);
			i = mask ? (i & mask) : (i < size ? i : i % size);
			#pragma unroll
			for (int f = 0; f < F; ++f) v[c][f] = __half2float(lvl[(size_t)i * F + f]);
		}
		const float wx[2] = {1.0f - fx, fx}, wy[2] = {1.0f - fy, fy}, wz[2] = {1.0f - fz, fz};
		#pragma unroll
		for (int f = 0; f < F; ++f) {
			const float ge = genc[l * F + f] * scale;
			float dx = 0.0f, dy = 0.0f, dz = 0.0f;
			#pragma unroll
			for (int b = 0; b < 2; ++b) {
				#pragma unroll
				for (int a2 = 0; a2 < 2; ++a2) {
					dx += wy[a2] * wz[b] * (v[1 + 2 * a2 + 4 * b][f] - v[0 + 2 * a2 + 4 * b][f]);
					dy += wx[a2] * wz[b] * (v[a2 + 2 + 4 * b][f] - v[a2 + 4 * b][f]);
					dz += wx[a2] * wy[b] * (v[a2 + 2 * b + 4][f] - v[a2 + 2 * b][f]);
				}
			}
			gx += ge * dx; gy += ge * dy; gz += ge * dz;
		}
	}
	// network_to_density_derivative (nerf_device.cuh:245-254), normal = -derivative * gradient, normalised
	float der = 1.0f;
	if (density_act == 1) der = raw > 0.0f ? 1.0f : 0.0f;
	else if (density_act == 2) { const float d = 1.0f / (1.0f + expf(-raw)); der = d * (1.0f - d); }
	else if (density_act == 3) der = expf(fminf(fmaxf(raw, -15.0f), 15.0f));
	float nx = -der * gx, ny = -der * gy, nz = -der * gz;
	const float len = sqrtf(nx * nx + ny * ny + nz * nz);
	nx /= len; ny /= len; nz /= len;
	uint2 o = out[idx];
	o.x = pack_h2(nx, ny);
	o.y = (o.y & 0xFFFF0000u) | (pack_h2(nz, 0.0f) & 0xFFFFu);
	out[idx] = o;
}
void launch_normals(const NetCfg& cfg, const GridLevels& g, const __half* params, const FrameArgs& fa, const float4* samples, const uint32_t* sample_counts,
	const uint32_t* n_chunks_ptr, uint32_t max_chunks, uint2* out, cudaStream_t s) {
	if (!max_chunks) return;
	const float3 lo = make_float3(fa.train_aabb[0], fa.train_aabb[1], fa.train_aabb[2]);
	const float3 diag = make_float3(fa.train_aabb[3] - fa.train_aabb[0], fa.train_aabb[4] - fa.train_aabb[1], fa.train_aabb[5] - fa.train_aabb[2]);
	const size_t smem = (size_t)(64 * 32 + (cfg.density_hidden - 1) * 64 * 64 + 16 * 64) * sizeof(__half);
	const uint32_t blocks = (max_chunks * CHUNK + 127) / 128;
	if (cfg.F == 4) normals_kernel<4><<<blocks, 128, smem, s>>>(g, params + cfg.grid_offset, params + cfg.density_w_offset, (int)cfg.density_hidden, lo, diag, samples, sample_counts, n_chunks_ptr, max_chunks, fa.density_act, out);
	else normals_kernel<2><<<blocks, 128, smem, s>>>(g, params + cfg.grid_offset, params + cfg.density_w_offset, (int)cfg.density_hidden, lo, diag, samples, sample_counts, n_chunks_ptr, max_chunks, fa.density_act, out);
}

void launch_eval(const NetCfg& cfg, const GridLevels& g, const __half* params, const FrameArgs& fa, const float4* samples,
	const uint32_t* sample_counts, const uint32_t* alive_list, RaySoA rays, const float* dirs_override, const uint32_t* n_chunks_ptr,
	uint32_t max_chunks, uint2* out, int impl, float4* fused_frame, float* fused_depth, uint32_t* survive, uint32_t* counters, cudaStream_t s,
	bool density_only, const ParamTextures* tex) {
	if (!max_chunks) return;
	EvalArgs a;
	a.density_only = (density_only && impl == 1) ? 1 : 0; // the warp-MMA kernel always runs both MLPs (sigma is the same)
	a.g = g;
	a.table = params + cfg.grid_offset;
	a.w_density = params + cfg.density_w_offset;
	a.w_rgb = params + cfg.rgb_w_offset;
	a.density_hidden = (int)cfg.density_hidden; a.rgb_hidden = (int)cfg.rgb_hidden;
	for (int i = 0; i < 3; ++i) { a.train_lo[i] = fa.train_aabb[i]; a.train_diag[i] = fa.train_aabb[3 + i] - fa.train_aabb[i]; }
	a.samples = samples; a.sample_counts = sample_counts; a.alive_list = alive_list; a.d_mw = rays.d_mw;
	a.dirs_override = dirs_override; a.n_chunks_ptr = n_chunks_ptr; a.n_chunks = max_chunks; a.out = out;
	a.fused = (impl == 1 && fused_frame) ? 1 : 0;
	a.unit_box = (a.train_lo[0] == 0.f && a.train_lo[1] == 0.f && a.train_lo[2] == 0.f && a.train_diag[0] == 1.f && a.train_diag[1] == 1.f && a.train_diag[2] == 1.f) ? 1 : 0;
	a.ray_sh = (impl == 1 && alive_list && !dirs_override) ? rays.sh : nullptr;
	a.gather_variant = fa.eval_gather; a.groups = fa.eval_groups; a.tex_split = fa.eval_texsplit;
	a.tex16 = 0; a.tex8 = 0; a.tex_base = (uint32_t)(cfg.grid_offset / 4);
	if (impl == 1 && tex && tex->tex16 && cfg.F == 4 && cfg.grid_offset % 8 == 0) { a.tex16 = tex->tex16; a.tex8 = tex->tex8; }
	a.comp = make_comp_args(fa, rays, alive_list, samples, fused_frame, fused_depth, survive, counters);
	if (impl == 1) { launch_eval_tc(a, cfg.F, max_chunks, s); return; }
	size_t smem = mlp_smem_bytes(a.density_hidden) + mlp_smem_bytes(a.rgb_hidden) + 4 * 32 * A_STRIDE;
	static int sm_count = 0;
	if (!sm_count) { int dev; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev); }
	uint32_t blocks = (max_chunks + 3) / 4;
	uint32_t persistent = (uint32_t)sm_count * 3u;
	if (blocks > persistent) blocks = persistent;
	if (cfg.F == 4) {
		cudaFuncSetAttribute(eval_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
		eval_kernel<4><<<blocks, 128, smem, s>>>(a);
	} else {
		cudaFuncSetAttribute(eval_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
		eval_kernel<2><<<blocks, 128, smem, s>>>(a);
	}
}

} // namespace hn
