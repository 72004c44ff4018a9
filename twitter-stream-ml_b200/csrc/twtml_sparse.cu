// placeholder until the CSR path lands (next commit)
#include "host.h"
namespace twtml {
int csr_launch_iter(twtml_ctx*, const twtml_model*, const twtml_batch*, int, double*, double*, int) { return fail(TWTML_E_INVALID_ARG, "CSR path not built"); }
int csr_partial_width(const twtml_batch* b) { return (int)b->d + 2; }
int csr_reduce(twtml_ctx*, const twtml_batch*, double*, const TrainState*) { return fail(TWTML_E_INVALID_ARG, "CSR path not built"); }
}
using namespace twtml;
extern "C" {
int32_t twtml_batch_from_tweets(twtml_ctx*, const uint16_t*, const int64_t*, const int32_t*, const int32_t*, const int32_t*, const int32_t*, const int64_t*, const uint8_t*, int64_t, int64_t, int32_t, int32_t, int32_t, uint32_t, twtml_batch**) { return fail(TWTML_E_INVALID_ARG, "not built"); }
int32_t twtml_batch_from_csr(twtml_ctx*, int64_t, int64_t, const int64_t*, const int32_t*, const double*, const double*, twtml_batch**) { return fail(TWTML_E_INVALID_ARG, "not built"); }
int32_t twtml_batch_synth_csr(twtml_ctx*, int64_t, int32_t, uint64_t, int64_t, twtml_batch**) { return fail(TWTML_E_INVALID_ARG, "not built"); }
int32_t twtml_batch_export_csr(const twtml_batch*, int64_t*, int32_t*, double*, double*) { return fail(TWTML_E_INVALID_ARG, "not built"); }
}
