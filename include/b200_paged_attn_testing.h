/*
 * b200_paged_attn_testing.h -- process-wide TEST HOOKS of libb200pa.so.
 *
 * Not part of the product ABI (include/b200_paged_attn.h): these exist so that the GPU test
 * suite can force one kernel family, the tuning tools can pick an alternative instantiation and
 * tools/ktrace.py can time the phases of the decode kernel.  They are compiled in only when the
 * library is built with -DB200PA_TESTING=1 (vllm_main_modify_b200/build.py does, unless
 * B200PA_TESTING=0 is in the environment); a build without them always behaves as
 * B200PA_IMPL_AUTO / tune 0 / no trace and exports none of these symbols.
 */
#ifndef B200_PAGED_ATTN_TESTING_H_
#define B200_PAGED_ATTN_TESTING_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* which kernel family b200pa_paged_attention_v1/v2 may use */
enum {
  B200PA_IMPL_AUTO = 0,    /* stream kernel when the shape is covered and a workspace is given,
                              else generic */
  B200PA_IMPL_GENERIC = 1, /* SIMT kernel, whole reference domain, no workspace */
  B200PA_IMPL_STREAM = 2   /* persistent stream-K tensor-core kernel (v1 and v2); error if the
                              shape is not covered or no workspace is given */
};

/* Override of the kernel family (default B200PA_IMPL_AUTO). */
void b200pa_set_impl(int impl);
int b200pa_get_impl(void);

/* Alternative (warps, stages, pages per stage, CTAs per SM) instantiation of the stream kernel
 * for head size 128 with a bf16 query (launch_stream in csrc/paged_attention_stream.cu); 0 = the
 * default.  The environment variable B200PA_STREAM_TUNE sets the initial value. */
void b200pa_testing_set_stream_tune(int tune);

/* Device buffer of `bytes` bytes (>= 256 per CTA of the decode kernel, 8-byte aligned) into which
 * every following stream-kernel launch writes %globaltimer stamps of its phases (slot layout:
 * StreamParams::trace); NULL switches tracing off.  The caller owns the buffer. */
void b200pa_testing_set_trace(void* device_buffer, int64_t bytes);

#ifdef __cplusplus
}
#endif
#endif /* B200_PAGED_ATTN_TESTING_H_ */
