// vv_dec.cu — placeholder entry points (replaced by the real decoder); they fail loudly.
#include "../../include/vvb200.h"
#include "vv_common.cuh"
#define UNSUP(name) do { vv_set_error(name ": not implemented yet"); return VV_ERR_UNSUPPORTED; } while (0)
extern "C" {
int vv_dec_create(const VVDecDesc*, const void* const*, int, VVDec**) { UNSUP("vv_dec_create"); }
void vv_dec_destroy(VVDec*) {}
int vv_dec_decode(VVDec*, const void*, int, int, int, float, float, void*, int, void*) { UNSUP("vv_dec_decode"); }
int vv_cache_create(VVDec*, int, VVCache**) { UNSUP("vv_cache_create"); }
void vv_cache_destroy(VVCache*) {}
int vv_cache_zero(VVCache*, const int64_t*, int, void*) { UNSUP("vv_cache_zero"); }
int vv_cache_zero_all(VVCache*, void*) { UNSUP("vv_cache_zero_all"); }
int vv_dec_step(VVDec*, VVCache*, const void*, const int64_t*, int, int, int, float, float, void*, int, void*) { UNSUP("vv_dec_step"); }
}
