#include "orb_internal.h"
extern "C" {
int orb_matcher_create(int, int, int, orb_matcher**) { orb_set_error("wip"); return ORB_ERR_INTERNAL; }
void orb_matcher_destroy(orb_matcher*) {}
void* orb_matcher_stream(const orb_matcher*) { return nullptr; }
long long orb_matcher_launch_count(const orb_matcher*) { return 0; }
int orb_matcher_sync(orb_matcher*) { return ORB_ERR_INTERNAL; }
int orb_descriptor_distance(orb_matcher*, const uint8_t*, const uint8_t*, int, int32_t*) { return ORB_ERR_INTERNAL; }
int orb_match_bruteforce(orb_matcher*, const uint8_t*, const uint8_t*, int, int, int, int32_t*, int32_t*, int32_t*) { return ORB_ERR_INTERNAL; }
int orb_match_bruteforce_device(orb_matcher*, const uint8_t*, const uint8_t*, int, int, int, int32_t*, int32_t*, int32_t*) { return ORB_ERR_INTERNAL; }
int orb_assign_features_to_grid(orb_matcher*, const float*, const float*, int, const orb_grid_params*, int32_t*, int32_t*, int32_t*) { return ORB_ERR_INTERNAL; }
int orb_search_by_projection_mappoints(orb_matcher*, const orb_frame_view*, const orb_mappoint_view*, float, float, const uint8_t*, int32_t*, int32_t*) { return ORB_ERR_INTERNAL; }
int orb_search_by_projection_lastframe(orb_matcher*, const orb_frame_view*, const orb_lastframe_view*, const orb_camera*, float, int, int, int, const uint8_t*, int32_t*, int32_t*) { return ORB_ERR_INTERNAL; }
int orb_search_by_bow(orb_matcher*, const orb_frame_view*, const orb_featvec_view*, const uint8_t*, const orb_frame_view*, const orb_featvec_view*, float, int, int32_t*, int32_t*) { return ORB_ERR_INTERNAL; }
int orb_search_for_triangulation(orb_matcher*, const orb_frame_view*, const orb_featvec_view*, const uint8_t*, const orb_frame_view*, const orb_featvec_view*, const uint8_t*, const float*, float, float, const float*, int, int, int32_t*, int32_t*) { return ORB_ERR_INTERNAL; }
}
