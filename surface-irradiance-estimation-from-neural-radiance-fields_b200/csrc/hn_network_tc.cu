// (b)+(c) Fused network evaluation on the 5th-generation tensor cores (tcgen05 / TMEM), sm_100a only.
//
// One CTA = 128 threads = one tile of 128 samples (4 ray chunks of 32; thread t owns sample row t = TMEM lane t).
//   gather   : thread t computes the 32 hash-grid features of its sample (tcnn grid.h:48-212 arithmetic, bit-identical to the
//              oracle) and stores them as the A operand in shared memory in the UMMA canonical K-major layout
//   layers   : ONE elected thread issues tcgen05.mma (M=128, N=64 or 16, K=16 per instruction, fp16 x fp16 -> fp32 in TMEM);
//              completion arrives on an mbarrier via tcgen05.commit
//   epilogue : each warp reads its 32 TMEM lanes with tcgen05.ld, applies ReLU, rounds to fp16 (tcnn stores fp16
//              activations between layers, fully_fused_mlp.cu:47-129) and writes the next layer's A operand
//   output   : (r,g,b,sigma) fp16, 8 B per sample, coalesced
// All weights (<= 28 KB) stay resident in shared memory in the canonical B layout for the life of the (persistent) CTA;
// the 64 B/sample features and every hidden activation never leave the SM. Several CTAs per SM overlap one tile's gather
// (L1/L2 latency bound) with another tile's MMA/epilogue rounds.
//
// Shared-memory operand layout (no swizzle, K-major; CUTLASS cute/atom/mma_traits_sm100.hpp "LayoutType::INTERLEAVE"):
//   a core matrix is 8 rows x 16 B, rows contiguous; here every 16-byte K-chunk c of the tile is one plane
//   [c][row][8 halfs], so element (row, k) sits at c*(ROWS*16) + row*16 + (k%8)*2 with c = k/8. For one K=16 instruction
//   the descriptor holds start = plane 2*kk, LBO (next K-chunk) = ROWS*16, SBO (next 8-row group) = 128.
#include "hn_network.cuh"
#include "hn_shade.cuh"
#include "hn_tc.cuh"
#include <map>
#include <mutex>
#include <utility>

namespace hn {
namespace tc {


// SH input of the colour MLP for one ray direction: warp_direction (d+1)/2 as the reference stores it in the NerfCoordinate, back to
// [-1,1] inside kernel_sh (spherical_harmonics.h:46-72), 16 coefficients rounded to fp16
__device__ __forceinline__ void dir_sh(float dx, float dy, float dz, uint4& s0, uint4& s1) {
	dx = __fmul_rn(__fadd_rn(dx, 1.0f), 0.5f) * 2.f - 1.f;
	dy = __fmul_rn(__fadd_rn(dy, 1.0f), 0.5f) * 2.f - 1.f;
	dz = __fmul_rn(__fadd_rn(dz, 1.0f), 0.5f) * 2.f - 1.f;
	float sh[16];
	sh4(dx, dy, dz, sh);
	s0 = make_uint4(pack_h2(sh[0], sh[1]), pack_h2(sh[2], sh[3]), pack_h2(sh[4], sh[5]), pack_h2(sh[6], sh[7]));
	s1 = make_uint4(pack_h2(sh[8], sh[9]), pack_h2(sh[10], sh[11]), pack_h2(sh[12], sh[13]), pack_h2(sh[14], sh[15]));
}

// The direction of a ray does not change over the passes of a frame: its SH coefficients are computed once, for the rays of the
// first live list, instead of by every one of the 32 lanes of every chunk of every pass (~60 instructions per sample).
__global__ void ray_sh_kernel(const uint32_t* __restrict__ list, const uint32_t* __restrict__ n_ptr, uint32_t n_max, const float4* __restrict__ d_mw, uint4* __restrict__ out) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	const uint32_t n = n_ptr ? min(*n_ptr, n_max) : n_max;
	if (i >= n) return;
	const uint32_t ray = list ? list[i] : i;
	const float4 d = __ldg(d_mw + ray);
	uint4 s0, s1;
	dir_sh(d.x, d.y, d.z, s0, s1);
	out[2 * (size_t)ray] = s0;
	out[2 * (size_t)ray + 1] = s1;
}


// GV (gather variant): 0 = encode_sample (any layout), 1 = encode_sample_pow2 with paired 16-byte fetches on hashed levels,
// 2 = encode_sample_pow2 with 8 independent 8-byte fetches on hashed levels. 1 and 2 need the base layout (launcher checks).
template <int F, int G, bool FUSED, bool DONLY, uint32_t TEXMASK = 0, int NDENSE = -1, int GV = 0>
__global__ void __launch_bounds__(TILE * G, G >= 4 ? 2 : (G == 3 ? 2 : 4)) eval_tc_kernel(const __grid_constant__ EvalArgs a) {
	extern __shared__ __align__(1024) uint8_t smem[];
	// warp index through a shuffle from lane 0: the compiler then knows it is warp-uniform and keeps everything derived from it
	// (tile group, TMEM / shared-memory addresses, descriptors of the MMA issue path) in uniform registers
	const int warp_u = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
	const int grp = warp_u >> 2;                          // tile group of this thread
	const int lane = threadIdx.x & 31, warp = warp_u & 3, tid = warp * 32 + lane; // within the group; warp == CTA warp index % 4
	uint8_t* a_s = smem + (size_t)grp * A_BYTES;          // this group's A operand, 8 planes x 128 rows x 16 B
	uint8_t* wd_s = smem + (size_t)G * A_BYTES;           // density MLP weights
	uint8_t* wr_s = wd_s + mlp_bytes(a.density_hidden);   // colour MLP weights
	uint64_t* bar_p = reinterpret_cast<uint64_t*>(wr_s + mlp_bytes(a.rgb_hidden)) + grp;
	uint32_t* tmem_p = reinterpret_cast<uint32_t*>(reinterpret_cast<uint64_t*>(wr_s + mlp_bytes(a.rgb_hidden)) + G);

	// ---- one-time set-up: weights -> canonical B planes, mbarrier, TMEM columns
	{
		uint8_t* d = wd_s;
		const __half* w = a.w_density;
		d += stage_layer(w, 64, 32, d); w += 64 * 32;
		for (int l = 1; l < a.density_hidden; ++l) { d += stage_layer(w, 64, 64, d); w += 64 * 64; }
		stage_layer(w, 16, 64, d);
		d = wr_s; w = a.w_rgb;
		d += stage_layer(w, 64, 32, d); w += 64 * 32;
		for (int l = 1; l < a.rgb_hidden; ++l) { d += stage_layer(w, 64, 64, d); w += 64 * 64; }
		stage_layer(w, 16, 64, d);
	}
	const uint32_t bar = smem_addr(bar_p);
	if (tid == 0) { mbar_init(bar, 1); fence_barrier_init(); }
	constexpr uint32_t TCOLS = G == 3 ? 256u : TMEM_COLS * G; // tcgen05.alloc takes a power of two >= 32
	if (threadIdx.x < 32) tmem_alloc(smem_addr(tmem_p), TCOLS);
	fence_proxy_async();   // the staged weights are read by the async proxy (tcgen05.mma)
	tc_fence_before();
	__syncthreads();
	tc_fence_after();
	const uint32_t tmem_base = *tmem_p + (uint32_t)grp * TMEM_COLS;
	const uint32_t t_row = tmem_base + ((uint32_t)(warp * 32) << 16); // this warp's 32 lanes, column 0
	const uint32_t a_addr = smem_addr(a_s), wd_addr = smem_addr(wd_s), wr_addr = smem_addr(wr_s);
	uint8_t* a_row = a_s + tid * 16;
	uint32_t phase = 0;
	uint32_t composited_acc = 0;

	const uint32_t n_chunks = a.n_chunks_ptr ? min(*a.n_chunks_ptr, a.n_chunks) : a.n_chunks;
	const uint32_t n_tiles = (n_chunks + 3) / 4;

	// every thread of the CTA runs the same number of rounds per tile: barriers are CTA-uniform by construction
	#define HN_ROUND(K, N, W_ADDR) do { \
		fence_proxy_async(); tc_fence_before(); group_sync(grp); \
		if (tid == 0) { tc_fence_after(); issue_layer(tmem_base, a_addr, (W_ADDR), (K), (N), bar); } \
	} while (0)
	#define HN_WAIT() do { mbar_wait(bar, phase); phase ^= 1u; tc_fence_after(); } while (0)

	for (uint32_t tile = blockIdx.x * G + grp; tile < n_tiles; tile += gridDim.x * G) {
		const uint32_t chunk = tile * 4 + warp;
		const uint32_t count = chunk < n_chunks ? (a.sample_counts ? a.sample_counts[chunk] : CHUNK) : 0u;
		// ---- (b) gather: 32 features of this thread's sample -> A planes 0..3
		if (GV != 0) {
			if ((uint32_t)lane < count && a.tex_split != 9) { // eval_texsplit 9 / 8: profiling switches - MLPs only (zero features) / gather only
				const float4 s = __ldg(a.samples + (size_t)chunk * CHUNK + lane);
				float px = s.x, py = s.y, pz = s.z;
				if (!a.unit_box) { // warp_position; (p - 0) / 1 is the identity, so the unit cube skips three IEEE divisions
					px = __fdiv_rn(__fsub_rn(s.x, a.train_lo[0]), a.train_diag[0]);
					py = __fdiv_rn(__fsub_rn(s.y, a.train_lo[1]), a.train_diag[1]);
					pz = __fdiv_rn(__fsub_rn(s.z, a.train_lo[2]), a.train_diag[2]);
				}
				encode_sample_pow2<TEXMASK, NDENSE, GV == 1>(a.g, a.table, px, py, pz, a_row, A_PLANE, a.tex16, a.tex8, a.tex_base);
			} else {
				#pragma unroll
				for (int c = 0; c < 4; ++c) *reinterpret_cast<uint4*>(a_row + c * A_PLANE) = make_uint4(0u, 0u, 0u, 0u);
			}
		} else {
			uint32_t enc[16];
			#pragma unroll
			for (int k = 0; k < 16; ++k) enc[k] = 0u;
			if ((uint32_t)lane < count) {
				const float4 s = __ldg(a.samples + (size_t)chunk * CHUNK + lane);
				float px = s.x, py = s.y, pz = s.z;
				if (!a.unit_box) { // warp_position; (p - 0) / 1 is the identity, so the unit cube skips three IEEE divisions
					px = __fdiv_rn(__fsub_rn(s.x, a.train_lo[0]), a.train_diag[0]);
					py = __fdiv_rn(__fsub_rn(s.y, a.train_lo[1]), a.train_diag[1]);
					pz = __fdiv_rn(__fsub_rn(s.z, a.train_lo[2]), a.train_diag[2]);
				}
				encode_sample<F, TEXMASK, NDENSE>(a.g, a.table, px, py, pz, enc, a.tex16, a.tex8, a.tex_base);
			}
			#pragma unroll
			for (int c = 0; c < 4; ++c)
				*reinterpret_cast<uint4*>(a_row + c * A_PLANE) = make_uint4(enc[4 * c], enc[4 * c + 1], enc[4 * c + 2], enc[4 * c + 3]);
		}
		if (GV != 0 && a.tex_split == 8) continue;
		// ---- (c) density MLP
		uint32_t w_addr = wd_addr;
		HN_ROUND(32, 64, w_addr); w_addr += 64 * 32 * 2;
		for (int l = 1; l < a.density_hidden; ++l) {
			HN_WAIT(); hidden_epilogue(t_row, a_row);
			HN_ROUND(64, 64, w_addr); w_addr += 64 * 64 * 2;
		}
		HN_WAIT(); hidden_epilogue(t_row, a_row);
		HN_ROUND(64, 16, w_addr);
		if (DONLY) { // light rays only need sigma: density channel 0, rounded to fp16 like the full path
			HN_WAIT();
			uint32_t v[4];
			tmem_ld4(t_row, v);
			tc_wait_ld();
			if (chunk < n_chunks) a.out[(size_t)chunk * CHUNK + lane] = make_uint2(0u, (pack_plain(v[0], 0u) & 0xFFFFu) << 16);
			tc_fence_before(); // the next tile's first MMA overwrites these TMEM columns
			continue;
		}
		// ---- colour MLP input: [16 density outputs | 16 SH coefficients of the ray direction]
		uint32_t sigma_bits;
		{
			uint4 s0, s1;
			if (a.ray_sh) { // computed once per ray and frame by ray_sh_kernel: one broadcast 32-byte read per warp
				const uint32_t ray = chunk < n_chunks ? a.alive_list[chunk] : 0u;
				s0 = __ldg(a.ray_sh + 2 * (size_t)ray);
				s1 = __ldg(a.ray_sh + 2 * (size_t)ray + 1);
			} else {
				float dx = 0.f, dy = 0.f, dz = 1.f;
				if (chunk < n_chunks) {
					if (a.dirs_override) { dx = a.dirs_override[3 * chunk]; dy = a.dirs_override[3 * chunk + 1]; dz = a.dirs_override[3 * chunk + 2]; }
					else { const float4 d = __ldg(a.d_mw + a.alive_list[chunk]); dx = d.x; dy = d.y; dz = d.z; }
				}
				dir_sh(dx, dy, dz, s0, s1);
			}
			HN_WAIT();
			uint32_t v[16];
			tmem_ld16(t_row, v);
			tc_wait_ld();
			const uint4 d0 = make_uint4(pack_plain(v[0], v[1]), pack_plain(v[2], v[3]), pack_plain(v[4], v[5]), pack_plain(v[6], v[7]));
			const uint4 d1 = make_uint4(pack_plain(v[8], v[9]), pack_plain(v[10], v[11]), pack_plain(v[12], v[13]), pack_plain(v[14], v[15]));
			sigma_bits = d0.x & 0xFFFFu; // density channel 0 (extract_density, nerf_network.h:31-43)
			*reinterpret_cast<uint4*>(a_row + 0 * A_PLANE) = d0;
			*reinterpret_cast<uint4*>(a_row + 1 * A_PLANE) = d1;
			*reinterpret_cast<uint4*>(a_row + 2 * A_PLANE) = s0;
			*reinterpret_cast<uint4*>(a_row + 3 * A_PLANE) = s1;
		}
		w_addr = wr_addr;
		HN_ROUND(32, 64, w_addr); w_addr += 64 * 32 * 2;
		for (int l = 1; l < a.rgb_hidden; ++l) {
			HN_WAIT(); hidden_epilogue(t_row, a_row);
			HN_ROUND(64, 64, w_addr); w_addr += 64 * 64 * 2;
		}
		HN_WAIT(); hidden_epilogue(t_row, a_row);
		HN_ROUND(64, 16, w_addr);
		// ---- output (r,g | b,sigma), 8 B per sample; fused: composite the chunk in this warp
		const bool fuse_here = FUSED && chunk < n_chunks;
		RayPre pre;
		if (fuse_here) pre = composite_prefetch(a.comp, chunk, count, lane);
		HN_WAIT();
		{
			uint32_t v[4];
			tmem_ld4(t_row, v);
			tc_wait_ld();
			uint2 o;
			o.x = pack_plain(v[0], v[1]);
			o.y = (pack_plain(v[2], 0u) & 0xFFFFu) | (sigma_bits << 16);
			if (!FUSED) {
				if (chunk < n_chunks) a.out[(size_t)chunk * CHUNK + lane] = o;
			} else if (fuse_here) {
				composited_acc += composite_chunk(a.comp, pre, chunk, count, lane, o);
			}
		}
	}
	if (FUSED) { // frame statistics: one atomic per warp per launch
		if (lane == 0 && composited_acc) atomicAdd((unsigned long long*)&a.comp.counters[CNT_COMPOSITED_LO], (unsigned long long)composited_acc);
	}
	#undef HN_ROUND
	#undef HN_WAIT
	tc_fence_before();
	__syncthreads();
	if (threadIdx.x < 32) tmem_dealloc(*tmem_p, TCOLS);
}

} // namespace tc

namespace {
// per-device launch state: SM count and "dynamic shared memory limit already raised for this kernel instance"
int sm_count_of_current_device() {
	static int counts[64] = {};
	int dev = 0;
	cudaGetDevice(&dev);
	if (dev < 0 || dev >= 64) { int n = 0; cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev); return n; }
	if (!counts[dev]) cudaDeviceGetAttribute(&counts[dev], cudaDevAttrMultiProcessorCount, dev);
	return counts[dev];
}
// cudaFuncAttributeMaxDynamicSharedMemorySize is raised once per (kernel instance, device), not on every launch. K is the same
// function-pointer TYPE for every instance, so the record is keyed by the pointer VALUE.
template <typename K>
void launch_tc(K kernel, const EvalArgs& a, uint32_t blocks, int threads, size_t smem, cudaStream_t s) {
	static std::mutex mu;
	static std::map<std::pair<const void*, int>, size_t> raised;
	int dev = 0;
	cudaGetDevice(&dev);
	{
		std::lock_guard<std::mutex> lock(mu);
		size_t& r = raised[std::make_pair((const void*)kernel, dev)];
		if (r < smem) {
			cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
			r = smem;
		}
	}
	kernel<<<blocks, threads, smem, s>>>(a);
}
} // namespace

void launch_ray_sh(const uint32_t* list, const uint32_t* n_ptr, uint32_t n_max, const float4* d_mw, uint4* out, cudaStream_t s) {
	if (!n_max) return;
	tc::ray_sh_kernel<<<(n_max + 255) / 256, 256, 0, s>>>(list, n_ptr, n_max, d_mw, out);
}

void launch_eval_tc(const EvalArgs& a, uint32_t F, uint32_t max_chunks, cudaStream_t s) {
	if (!max_chunks) return;
	constexpr int G = 4; // tile groups per CTA: 512 threads, 2 CTAs per SM = 8 tiles (32 warps) in flight per SM
	const int sm_count = sm_count_of_current_device();
	const uint32_t tiles = (max_chunks + 3) / 4;
	const uint32_t persistent = (uint32_t)sm_count * 2u;
	auto smem_for = [&](int g) { return (size_t)g * tc::A_BYTES + tc::mlp_bytes(a.density_hidden) + tc::mlp_bytes(a.rgb_hidden) + 8 * g + 16; };
	auto blocks_for = [&](int g) { const uint32_t ctas = (tiles + g - 1) / g; return ctas < persistent ? ctas : persistent; };
	const size_t smem = smem_for(G);
	const uint32_t blocks = blocks_for(G);
	#define HN_LAUNCH_TC(FF, FU, DO, TM, ND, GVV) launch_tc(tc::eval_tc_kernel<FF, G, FU, DO, TM, ND, GVV>, a, blocks, tc::TILE * G, smem, s)
	#define HN_LAUNCH_TC3(FF, FU, DO, TM, ND, GVV) launch_tc(tc::eval_tc_kernel<FF, 3, FU, DO, TM, ND, GVV>, a, blocks_for(3), tc::TILE * 3, smem_for(3), s)
	// F = 4 with texture objects: the five fine levels (3..7, the hashed ones of the base config) are gathered through the TEX
	// pipe, the three coarse ones through the LSU pipe. Measured on B200 (profiles/r01_tex_probe.txt): eval 7.03 -> 6.53 ms per
	// 1080p frame, bit-identical output; every other split of (level, corner pair) between the two pipes was slower.
	constexpr uint32_t TM = 0xFFFFF000u;
	// level layout of configs/nerf/base.json: 3 dense levels with even resolutions, then hashed levels, EVERY table a power of two
	// -> level kind compiled in and the mask-only gather (encode_sample_pow2)
	bool base_layout = F == 4 && a.g.L == 8;
	for (int l = 0; l < 8 && base_layout; ++l)
		base_layout = a.g.mask[l] != 0 && (l < 3 ? (a.g.hashed[l] == 0 && (a.g.res[l] & 1u) == 0 && (uint64_t)a.g.res[l] * a.g.res[l] * a.g.res[l] == a.g.size[l]) : a.g.hashed[l] != 0);
	const int gv = a.gather_variant;
	if (F == 4 && a.tex16 && base_layout && !a.fused) {
		// which (level, x-pair) fetches ride the TEX pipe (bit 4 l + p), the rest the LSU pipe: option eval_texsplit, picked from a sweep
		// on B200 (profiles/r02_eval_variants.txt) - with 8 independent 8-byte fetches per hashed level the TEX input queue throttles
		// when all five fine levels use it, so part of them goes back to LSU
		constexpr uint32_t TM1 = 0xFFF00000u; // levels 5..7
		constexpr uint32_t TM2 = 0x33333000u; // two of the four x-pairs of levels 3..7
		constexpr uint32_t TM3 = 0xFF000000u; // levels 6..7
		constexpr uint32_t TM4 = 0x0F0F0000u; // levels 4 and 6
		if (a.density_only) { if (gv == 2) HN_LAUNCH_TC(4, false, true, TM, 3, 2); else if (gv == 0) HN_LAUNCH_TC(4, false, true, TM, 3, 0); else HN_LAUNCH_TC(4, false, true, TM, 3, 1); }
		else if (a.groups == 3) { if (gv == 2) HN_LAUNCH_TC3(4, false, false, TM, 3, 2); else HN_LAUNCH_TC3(4, false, false, TM, 3, 1); }
		else if (gv == 2 && a.tex_split == 1) HN_LAUNCH_TC(4, false, false, TM1, 3, 2);
		else if (gv == 2 && a.tex_split == 2) HN_LAUNCH_TC(4, false, false, TM2, 3, 2);
		else if (gv == 2 && a.tex_split == 3) HN_LAUNCH_TC(4, false, false, TM3, 3, 2);
		else if (gv == 2 && a.tex_split == 4) HN_LAUNCH_TC(4, false, false, TM4, 3, 2);
		else if (gv == 2 && a.tex_split == 5) HN_LAUNCH_TC(4, false, false, 0u, 3, 2);
		else { if (gv == 2) HN_LAUNCH_TC(4, false, false, TM, 3, 2); else if (gv == 0) HN_LAUNCH_TC(4, false, false, TM, 3, 0); else HN_LAUNCH_TC(4, false, false, TM, 3, 1); }
	}
	else if (F == 4 && a.tex16) { if (a.density_only) HN_LAUNCH_TC(4, false, true, TM, -1, 0); else if (a.fused) HN_LAUNCH_TC(4, true, false, TM, -1, 0); else HN_LAUNCH_TC(4, false, false, TM, -1, 0); }
	else if (F == 4) { if (a.density_only) HN_LAUNCH_TC(4, false, true, 0u, -1, 0); else if (a.fused) HN_LAUNCH_TC(4, true, false, 0u, -1, 0); else HN_LAUNCH_TC(4, false, false, 0u, -1, 0); }
	else { if (a.density_only) HN_LAUNCH_TC(2, false, true, 0u, -1, 0); else if (a.fused) HN_LAUNCH_TC(2, true, false, 0u, -1, 0); else HN_LAUNCH_TC(2, false, false, 0u, -1, 0); }
	#undef HN_LAUNCH_TC
	#undef HN_LAUNCH_TC3
}

} // namespace hn
