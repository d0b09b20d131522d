/* Compiled as C by tests/test_host.py: include/gsb200.h must be a valid C header and every entry
 * point it declares must link against libgsb200.so. */
#include "gsb200.h"
#include <stdio.h>

int main(void) {
    void* syms[] = {
        (void*)gsb_version, (void*)gsb_last_error, (void*)gsb_create, (void*)gsb_destroy, (void*)gsb_dims_create,
        (void*)gsb_fill_domain, (void*)gsb_init_fields, (void*)gsb_exchange, (void*)gsb_calculate, (void*)gsb_swap,
        (void*)gsb_step, (void*)gsb_step_group, (void*)gsb_get_fields, (void*)gsb_get_fields_async,
        (void*)gsb_output_wait, (void*)gsb_get_ghosted, (void*)gsb_set_ghosted, (void*)gsb_alloc_host,
        (void*)gsb_free_host, (void*)gsb_sync, (void*)gsb_get_step, (void*)gsb_set_step, (void*)gsb_field_stats, (void*)gsb_pdf,
        (void*)gsb_get_unique_id, (void*)gsb_comm_init, (void*)gsb_comm_init_local, (void*)gsb_set_kernel,
        (void*)gsb_get_kernel, (void*)gsb_timers_enable, (void*)gsb_timers, (void*)gsb_halo_bytes,
        (void*)gsb_launch_count, (void*)gsb_compute_stream, (void*)gsb_calculate_arrays,
        (void*)gsb_field_stats_noalloc, (void*)gsb_halo_timers, (void*)gsb_device_count,
    };
    gsb_config cfg;
    int32_t dims[3];
    if (gsb_dims_create(8, dims) || dims[0] != 2 || dims[1] != 2 || dims[2] != 2) return 1;
    if (gsb_fill_domain(&cfg, 64, 8, 5) || cfg.coords[0] != 1 || cfg.coords[1] != 0 || cfg.coords[2] != 1) return 2;
    if (cfg.neighbors[GSB_WEST] != 1 || cfg.neighbors[GSB_EAST] != -1 || cfg.neighbors[GSB_UP] != 7) return 3;
    if (!gsb_fill_domain(&cfg, 10, 3, 0)) return 4; /* L must divide by dims */
    printf("%d symbols, version %d, sizeof(gsb_config) = %zu\n", (int)(sizeof syms / sizeof syms[0]), gsb_version(),
           sizeof(gsb_config));
    return 0;
}
