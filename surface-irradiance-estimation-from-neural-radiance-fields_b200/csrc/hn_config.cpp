// Network configuration: configs/nerf/*.json / configs/geometry/*.json (with "parent" merge) -> NetCfg + level table.
// Mirrors Testbed::reset_network (src/testbed.cu:3844-4212) for the NeRF network and tcnn's GridEncoding constructor
// (dependencies/tiny-cuda-nn/include/tiny-cuda-nn/encodings/grid.h:668-730) for the parameter layout.
#include "hn_context.h"
#include "hn_json.h"
#include <cmath>
#include <fstream>
#include <sstream>
#include <algorithm>
#include <limits>

namespace hn {

std::string read_text_file(const std::string& path) {
	std::ifstream f(path, std::ios::binary);
	if (!f) throw std::runtime_error("cannot open '" + path + "'");
	std::stringstream ss;
	ss << f.rdbuf();
	return ss.str();
}

static std::string dir_of(const std::string& path) {
	size_t p = path.find_last_of('/');
	return p == std::string::npos ? std::string(".") : path.substr(0, p);
}

// merge_parent_network_config (src/testbed.cu:86-97)
static Json merge_parent(const Json& child, const std::string& child_dir) {
	if (!child.contains("parent")) return child;
	std::string parent_path = child_dir + "/" + child.at("parent").str;
	std::string text = read_text_file(parent_path);
	Json parent = JsonParser(text).parse();
	parent = merge_parent(parent, dir_of(parent_path));
	json_merge_patch(parent, child);
	return parent;
}

static std::string lower(std::string s) { std::transform(s.begin(), s.end(), s.begin(), [](unsigned char c) { return (char)std::tolower(c); }); return s; }

// The reference evaluates grid_scale / grid_resolution (tcnn common_device.h:709-718) INSIDE kernel_grid, compiled with
// --use_fast_math: exp2f becomes MUFU.EX2 (ex2.approx.ftz) and `* base_resolution - 1` one FFMA (SASS of the reference's
// kernel_grid<__half,3,2>: FMUL.FTZ, MUFU.EX2, FFMA.FTZ ... -1, F2I.CEIL). Whenever per_level_scale is not a power of two
// (configs/geometry/base.json: b = 1.38) that differs from the host's exp2f in the last bits, which moves every sample
// inside its cell. The per-level scale / resolution the gather uses are therefore computed on the device with the same
// three instructions; the OFFSET table stays host arithmetic, as in tcnn's GridEncoding constructor (grid.h:692-721).
__global__ void level_scale_kernel(uint32_t L, float log2_per_level_scale, uint32_t base_res, float* scales, uint32_t* res) {
	const uint32_t level = threadIdx.x;
	if (level >= L) return;
	float e;
	asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(__fmul_rn((float)level, log2_per_level_scale)));
	const float scale = __fmaf_rn(e, (float)base_res, -1.0f);
	scales[level] = scale;
	res[level] = (uint32_t)ceilf(scale) + 1u;
}

static uint32_t mlp_params(uint32_t in, uint32_t hidden, uint32_t out_padded) { return 64u * in + (hidden - 1) * 64u * 64u + out_padded * 64u; }

void finalize_network(hn_context* ctx) {
	NetCfg& n = ctx->net;
	// per_level_scale: src/testbed.cu:3958-3966 (float exp/log), desired finest resolution 2048 * aabb_scale
	if (n.per_level_scale <= 0.0f && n.L > 1) {
		n.per_level_scale = std::exp(std::log(2048.0f * (float)ctx->aabb_scale / (float)n.base_res) / (float)(n.L - 1));
	}
	// offset table: grid.h:692-721
	const float log2_scale = std::log2(n.per_level_scale);
	uint32_t offset = 0;
	for (uint32_t i = 0; i < n.L; ++i) {
		const float scale = std::exp2f((float)i * log2_scale) * (float)n.base_res - 1.0f; // grid_scale, common_device.h:709-714
		const uint32_t resolution = (uint32_t)std::ceil(scale) + 1;                       // grid_resolution, :716-718
		const uint32_t max_params = std::numeric_limits<uint32_t>::max() / 2;
		uint32_t params_in_level = std::pow((float)resolution, 3) > (float)max_params ? max_params : resolution * resolution * resolution;
		params_in_level = (params_in_level + 7u) / 8u * 8u;
		params_in_level = std::min(params_in_level, 1u << n.log2_T);
		n.offsets[i] = offset;
		n.scales[i] = scale;
		n.res[i] = resolution;
		offset += params_in_level;
	}
	n.offsets[n.L] = offset;
	if (ctx->stream) { // device-side scale / resolution per level (see level_scale_kernel); a context without a GPU only answers layout queries
		HN_CUDA(cudaSetDevice(ctx->device));
		DevBuf<float> d_scales; DevBuf<uint32_t> d_res;
		d_scales.ensure(MAX_LEVELS); d_res.ensure(MAX_LEVELS);
		level_scale_kernel<<<1, 32>>>(n.L, log2_scale, n.base_res, d_scales.p, d_res.p);
		HN_CUDA(cudaGetLastError());
		HN_CUDA(cudaMemcpy(n.scales, d_scales.p, n.L * sizeof(float), cudaMemcpyDeviceToHost));
		HN_CUDA(cudaMemcpy(n.res, d_res.p, n.L * sizeof(uint32_t), cudaMemcpyDeviceToHost));
	}
	// parameter blob: [density MLP | rgb MLP | grid] (nerf_network.h:356-371)
	n.density_w_offset = 0;
	n.rgb_w_offset = mlp_params(n.L * n.F, n.density_hidden, 16);
	n.grid_offset = n.rgb_w_offset + mlp_params(32, n.rgb_hidden, 16);
	n.n_params = n.grid_offset + (size_t)offset * n.F;

	GridLevels& g = ctx->levels;
	g.L = n.L;
	for (uint32_t i = 0; i < n.L; ++i) {
		g.offset[i] = n.offsets[i];
		g.size[i] = n.offsets[i + 1] - n.offsets[i];
		g.res[i] = n.res[i];
		g.scale[i] = n.scales[i];
		// grid_index (common_device.h:690-707): the dense stride loop runs while stride <= hashmap_size
		uint32_t stride = 1;
		for (uint32_t dim = 0; dim < 3 && stride <= g.size[i]; ++dim) stride *= n.res[i];
		g.hashed[i] = g.size[i] < stride ? 1u : 0u;
		g.mask[i] = (g.size[i] & (g.size[i] - 1)) == 0 ? g.size[i] - 1 : 0u;
	}
	ctx->has_net = true;
	ctx->param_tex.ptr = nullptr; // texture views are rebuilt for the new layout on the next frame
	ctx->has_params = false;
	// aabb_scale drives the cascades and the cone angle (src/testbed_nerf.cu:2728-2740)
	int max_cascade = 0;
	while ((1 << max_cascade) < ctx->aabb_scale) ++max_cascade;
	ctx->opt.max_cascade = max_cascade;
	ctx->opt.cone_angle_constant = ctx->aabb_scale <= 1 ? 0.0f : (1.0f / 256.0f);
}

void load_network_config_text(hn_context* ctx, const std::string& text, const std::string& base_dir) {
	Json cfg = JsonParser(text).parse();
	cfg = merge_parent(cfg, base_dir);
	apply_network_config(ctx, cfg);
}

// cfg: merged network config (also the root object of a snapshot; its "snapshot" member is ignored here)
void apply_network_config(hn_context* ctx, const Json& cfg) {
	if (!cfg.contains("encoding") || !cfg.contains("network")) throw std::runtime_error("network config needs 'encoding' and 'network' objects");
	const Json& enc = cfg.at("encoding");
	std::string otype = lower(enc.value("otype", std::string("OneBlob")));
	if (otype.find("grid") == std::string::npos) throw std::runtime_error("only (Hash)Grid position encodings are on this path, got '" + otype + "'");
	if (otype == "densegrid" || otype == "tiledgrid") throw std::runtime_error("only the Hash grid type is supported on this path");
	NetCfg n;
	n.F = (uint32_t)enc.value("n_features_per_level", 2.0);
	if (enc.contains("n_features") && enc.at("n_features").num > 0) n.L = (uint32_t)enc.at("n_features").num / n.F;
	else n.L = (uint32_t)enc.value("n_levels", 16.0);
	n.log2_T = (uint32_t)enc.value("log2_hashmap_size", 15.0);
	n.base_res = (uint32_t)enc.value("base_resolution", 0.0);
	if (!n.base_res) n.base_res = 1u << (n.log2_T / 3);
	n.per_level_scale = (float)enc.value("per_level_scale", 0.0);
	if (!(n.F == 2 || n.F == 4) || n.L * n.F != 32) throw std::runtime_error("this build supports 32 grid features: (n_levels, n_features_per_level) = (8,4) or (16,2)");
	if (n.L > MAX_LEVELS) throw std::runtime_error("too many grid levels");
	const std::string interp = lower(enc.value("interpolation", std::string("linear")));
	if (interp != "linear") throw std::runtime_error("only linear grid interpolation is supported");

	auto read_mlp = [&](const Json& j, const char* what) -> uint32_t {
		std::string t = lower(j.value("otype", std::string("FullyFusedMLP")));
		if (t != "fullyfusedmlp" && t != "cutlassmlp" && t != "megakernelmlp") throw std::runtime_error(std::string(what) + ": unsupported otype");
		if ((int)j.value("n_neurons", 64.0) != 64) throw std::runtime_error(std::string(what) + ": only n_neurons = 64 is supported");
		if (lower(j.value("activation", std::string("ReLU"))) != "relu") throw std::runtime_error(std::string(what) + ": only ReLU hidden activation is supported");
		if (lower(j.value("output_activation", std::string("None"))) != "none") throw std::runtime_error(std::string(what) + ": only output_activation None is supported");
		int h = (int)j.value("n_hidden_layers", 2.0);
		if (h < 1 || h > 8) throw std::runtime_error(std::string(what) + ": n_hidden_layers out of range");
		return (uint32_t)h;
	};
	n.density_hidden = read_mlp(cfg.at("network"), "network");
	n.rgb_hidden = cfg.contains("rgb_network") ? read_mlp(cfg.at("rgb_network"), "rgb_network") : 2u;
	if (cfg.contains("dir_encoding")) {
		const Json& de = cfg.at("dir_encoding");
		const Json* sh = &de;
		if (lower(de.value("otype", std::string(""))) == "composite" && de.contains("nested") && !de.at("nested").arr.empty()) sh = &de.at("nested").arr[0];
		if (lower(sh->value("otype", std::string("SphericalHarmonics"))) != "sphericalharmonics" || (int)sh->value("degree", 4.0) != 4)
			throw std::runtime_error("only SphericalHarmonics degree 4 direction encoding is supported");
	}
	ctx->net = n;
	{ // loss + optimizer stack (Ema > ExponentialDecay > Adam in configs/nerf/base.json:2-22); unknown wrappers are skipped
		TrainCfg t;
		if (cfg.contains("loss")) {
			const std::string lt = lower(cfg.at("loss").value("otype", std::string("Huber")));
			// ELossType (nerf_device.cuh:48-56, string_to_loss_type src/testbed_nerf.cu): ids used across the ABI
			static const char* names[] = {"l2", "huber", "relativel2", "l1", "mape", "smape", "logl1"};
			t.loss_type = -1;
			for (int k = 0; k < 7; ++k) if (lt == names[k]) t.loss_type = k;
			if (t.loss_type < 0) throw std::runtime_error("loss: unknown otype '" + lt + "'");
		}
		t.ema = false; t.exp_decay = false;
		const Json* o = cfg.find("optimizer");
		while (o && o->is_object()) {
			const std::string ot = lower(o->value("otype", std::string("Adam")));
			if (ot == "ema") { t.ema = true; t.ema_decay = (float)o->value("decay", 0.99); }
			else if (ot == "exponentialdecay") {
				t.exp_decay = true; t.decay_start = (uint32_t)o->value("decay_start", 0.0); t.decay_interval = (uint32_t)o->value("decay_interval", 10000.0);
				t.decay_base = (float)o->value("decay_base", 0.33);
			} else if (ot == "adam") {
				t.lr = (float)o->value("learning_rate", 1e-3); t.beta1 = (float)o->value("beta1", 0.9); t.beta2 = (float)o->value("beta2", 0.999);
				t.epsilon = (float)o->value("epsilon", 1e-8); t.l2_reg = (float)o->value("l2_reg", 1e-8);
			}
			o = o->find("nested");
		}
		ctx->train_cfg = t;
		ctx->train.ready = false;
	}
	{ // kept (without any snapshot payload) so hn_set_aabb_scale can re-derive per_level_scale, and for hn_save_snapshot
		Json keep; keep.type = Json::Object;
		for (auto& kv : cfg.obj) if (kv.first != "snapshot") keep.obj.push_back(kv);
		ctx->config_blob = msgpack_write(keep);
	}
	finalize_network(ctx);
}

void reapply_network_config(hn_context* ctx) { apply_network_config(ctx, msgpack_parse(ctx->config_blob)); }

} // namespace hn
