// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_common.h).
#pragma once
#include "oracle_common.h"
#include "oracle.h"

namespace orc {

struct GridLayout {
	uint32_t n_levels, F;
	float scale[32];
	uint32_t resolution[32];
	uint32_t offset[33];
};

struct NetworkView {
	const half_t *density_w1, *density_w2, *rgb_w1, *rgb_w2, *rgb_w3, *grid;
	GridLayout grid_layout;
};

GridLayout make_grid_layout(uint32_t n_levels, uint32_t F, uint32_t log2_hashmap_size, uint32_t base_resolution, float per_level_scale);
uint32_t grid_index(uint32_t hashmap_size, uint32_t resolution, uint32_t x, uint32_t y, uint32_t z);
void grid_encode(const GridLayout& g, const half_t* grid, vec3 x, half_t* out, uint32_t* out_idx);
void sh4_encode(vec3 dir_warped, half_t* out16);
void network_query(const NetworkView& net, vec3 pos_warped, vec3 dir_warped, half_t out4[4], half_t* feat32_out);
NetworkView make_network_view(const orc_model& m);

} // namespace orc
