// beam.cu — GPU beam search (decode_sequence_beam, seq2seq.py:1356-1544).  Under construction.
#include "model.h"
using namespace cor;
extern "C" {
int cor_beam_workspace_bytes(const cor_model*, int, int, int, const cor_beam_params*, size_t, size_t*) {
    set_error("beam search is not built yet");
    return 1;
}
int cor_correct_beam_host(cor_model*, const cor_lines*, const cor_beam_params*, void*, size_t,
                          int32_t*, float*, double*, int32_t*, float*, int64_t*, void*) {
    set_error("beam search is not built yet");
    return 1;
}
}
