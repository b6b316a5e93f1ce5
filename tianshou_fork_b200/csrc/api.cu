// Version / error plumbing of the C-ABI (include/ts_b200.h).
#include "ts_common.cuh"

namespace ts {
thread_local int32_t g_last_cuda_error = 0;
int g_tune_gae_rows_variant = 0;
int g_tune_gae_flat_variant = 0;
int g_tune_gae_flat_ctas = 0;
int g_tune_gae_flat_stages = 0;
int g_tune_segtree_group = 128;
int g_tune_segtree_ranged = 1;
int g_tune_pdl = 1;
int g_tune_gae_balance = 0;
int g_tune_gae_rows_packed = 1;
int g_tune_per_cluster = 0;
int g_tune_segtree_max_ctas = 148;
}

#include <string.h>

extern "C" int32_t ts_version(void) { return TS_B200_VERSION; }

extern "C" int32_t ts_last_cuda_error(void) { return ts::g_last_cuda_error; }

extern "C" const char* ts_error_string(int32_t code) {
    switch (code) {
        case TS_OK: return "ok";
        case TS_ERR_BAD_ARG: return "bad argument (null pointer, negative size or unknown dtype)";
        case TS_ERR_WORKSPACE: return "workspace too small";
        case TS_ERR_CUDA: return cudaGetErrorString((cudaError_t)ts::g_last_cuda_error);
        case TS_ERR_ALIGN: return "pointer not aligned to its element size";
        case TS_ERR_TOO_LARGE:
            return "size exceeds the fixed scratch capacity of this *_static entry point";
        default: return "unknown error code";
    }
}

extern "C" int32_t ts_tune_set(const char* key, int64_t value) {
    if (!key) return TS_ERR_BAD_ARG;
    if (strcmp(key, "gae_rows_variant") == 0) { ts::g_tune_gae_rows_variant = (int)value; return TS_OK; }
    if (strcmp(key, "gae_flat_variant") == 0 && value >= 0 && value <= 2) {
        ts::g_tune_gae_flat_variant = (int)value;
        return TS_OK;
    }
    if (strcmp(key, "gae_flat_ctas") == 0 && value >= 0 && value <= 3) {
        ts::g_tune_gae_flat_ctas = (int)value;
        return TS_OK;
    }
    if (strcmp(key, "gae_flat_stages") == 0 && value >= 0 && value <= 6) {
        ts::g_tune_gae_flat_stages = (int)value;
        return TS_OK;
    }
    if (strcmp(key, "segtree_max_ctas") == 0 && value >= 1 && value <= ts::kNumSMs) {
        ts::g_tune_segtree_max_ctas = (int)value;
        return TS_OK;
    }
    if (strcmp(key, "per_cluster") == 0 && (value == 0 || value == 8 || value == 16)) {
        ts::g_tune_per_cluster = (int)value;
        return TS_OK;
    }
    if (strcmp(key, "segtree_ranged") == 0 && (value == 0 || value == 1)) {
        ts::g_tune_segtree_ranged = (int)value;
        return TS_OK;
    }
    if (strcmp(key, "gae_rows_packed") == 0 && (value == 0 || value == 1)) {
        ts::g_tune_gae_rows_packed = (int)value;
        return TS_OK;
    }
    if (strcmp(key, "gae_balance") == 0 && (value == 0 || value == 1)) {
        ts::g_tune_gae_balance = (int)value;
        return TS_OK;
    }
    if (strcmp(key, "pdl") == 0 && (value == 0 || value == 1)) {
        ts::g_tune_pdl = (int)value;
        return TS_OK;
    }
    if (strcmp(key, "segtree_group") == 0 && (value == 128 || value == 256)) {
        ts::g_tune_segtree_group = (int)value;
        return TS_OK;
    }
    return TS_ERR_BAD_ARG;
}
