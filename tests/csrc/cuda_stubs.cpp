// CUDA-side entry points the host C++ refers to, stubbed for a host-only sanitizer build (tests only).
#include "gpualloc.h"
extern "C" {
int gpa_besteffort(gpa_ctx *, const int32_t *, int, const int32_t *, int, int, int, int32_t *, int *) { return GPA_ERR_CUDA; }
int gpa_subset_besteffort(gpa_ctx *, const int32_t *, int, const int32_t *, int, int, int, int32_t *, int *) { return GPA_ERR_CUDA; }
int gpa_get_topology(gpa_ctx *, int *, int32_t *, uint8_t *, int) { return GPA_ERR_CUDA; }
int gpa_set_device_order(gpa_ctx *, const int32_t *, int) { return GPA_ERR_CUDA; }
int gpa_set_topology_links(gpa_ctx *, int, const int32_t *) { return GPA_ERR_CUDA; }
const char *gpa_last_error(gpa_ctx *) { return ""; }
const char *gpa_strerror(int) { return "stub"; }
}
