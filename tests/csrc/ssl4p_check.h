/* ssl4p_check.h - C-ABI of tests/libssl4p_check.so: TEST INFRASTRUCTURE ONLY (an independent on-device implementation the
 * product kernels are cross-checked against; the product library libssl4p.so does not contain or call it).
 * Same contract as ssl4p_attention_fwd / ssl4p_attention_bwd of include/ssl4p.h, on warp-level mma.sync, T <= 208. */
#ifndef SSL4P_CHECK_H_
#define SSL4P_CHECK_H_
#ifdef __cplusplus
extern "C" {
#endif
const char* ssl4p_check_last_error(void);
int ssl4p_attention_fwd_mma(const void* qkv, void* out, float* lse, int B, int T, int H, int hd, float scale, void* stream);
int ssl4p_attention_bwd_mma(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv,
                            int B, int T, int H, int hd, float scale, void* stream);
#ifdef __cplusplus
}
#endif
#endif
