// temporary: replaced by sort.cu / bh.cu / sh.cu
#include "common.cuh"
namespace nb { void bh_destroy(nb_context*) {} void sh_destroy(nb_context*) {} }
#define NI { nb::set_error("not implemented"); return NB_ERR_INTERNAL; }
extern "C" {
int nb_bh_build(nb_context*, const nb_particles*, nb_stream) NI
int nb_bh_forces(nb_context*, const nb_particles*, float, float, float, nb_stream) NI
int nb_bh_export(nb_context*, uint32_t*, int32_t*, int32_t*, int32_t*, int32_t*, int32_t*, float*, float*) NI
int nb_bh_stats(nb_context*, uint64_t*, int*) NI
int nb_sh_build(nb_context*, const nb_particles*, float, nb_stream) NI
int nb_sh_forces(nb_context*, const nb_particles*, float, float, float, nb_stream) NI
int nb_sh_export(nb_context*, int32_t*, float*, int32_t*, int32_t*, int32_t*, int32_t*) NI
int nb_sh_total_cells(nb_context*, int*) NI
int nb_sort_pairs(nb_context*, uint32_t*, int32_t*, size_t, int, nb_stream) NI
int nb_bounding_box(nb_context*, const nb_particles*, float*, nb_stream) NI
}
