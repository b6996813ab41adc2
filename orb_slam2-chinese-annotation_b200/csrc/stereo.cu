#include "orb_internal.h"
extern "C" {
int orb_stereo_match(orb_extractor*, orb_extractor*, int, float, float, float, float*, float*, int, int32_t*) { orb_set_error("wip"); return ORB_ERR_INTERNAL; }
int orb_stereo_match_batch(orb_extractor*, orb_extractor*, int, float, float, float, float*, float*, int, int32_t*) { orb_set_error("wip"); return ORB_ERR_INTERNAL; }
int orb_stereo_match_device(orb_extractor*, orb_extractor*, int, float, float, float) { orb_set_error("wip"); return ORB_ERR_INTERNAL; }
int orb_stereo_device_results(const orb_extractor*, const float**, const float**) { orb_set_error("wip"); return ORB_ERR_INTERNAL; }
}
