// orbx_match.cu — placeholder while the extractor is brought up (replaced by the matcher kernels)
#include "orbx_internal.h"
#define STUB { orbx_set_error("not implemented yet"); return ORBX_E_STATE; }
int orbx_build_grids(orbx_handle*, int, int, int, int) STUB
extern "C" {
int orbx_search_init_pairs(orbx_handle*, int, int, int, float, int, int32_t*, int32_t*, int, uint64_t*) STUB
int orbx_extract_stereo_batch(orbx_handle*, const uint8_t*, const uint8_t*, size_t, size_t, int, int, int, int, int, float, float, float*, float*, int, uint64_t*) STUB
int orbm_descriptor_distance(const uint8_t*, const uint8_t*, int, int32_t*) STUB
int orbm_search_for_initialization(const orbm_frame*, const orbm_frame*, float*, int32_t*, int, float, int, int32_t*) STUB
int orbm_search_by_projection(int, const orbm_frame*, const orbm_proj_query*, const uint8_t*, int, const uint8_t*, uint8_t*, int32_t*, int32_t*, int32_t*, int, float, int, int32_t*) STUB
int orbm_compute_stereo_matches(const orbm_stereo_input*, float*, float*, uint64_t*) STUB
int orbm_knn2_device(const uint8_t*, int64_t, const uint8_t*, int64_t, int32_t, uint16_t*, int32_t*, void*) STUB
int orbm_knn2_merge_device(const uint16_t*, const int32_t*, int, int64_t, uint16_t*, int32_t*, void*) STUB
}
