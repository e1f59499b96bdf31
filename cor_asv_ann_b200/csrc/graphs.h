// graphs.h — executable CUDA graphs of the two decoder loops (beam search iterations, fast-greedy timesteps).
//
// Both loops are a WHILE conditional node whose body is one iteration (a handful of kernel launches) and whose
// condition a kernel of the body sets on the device: no launch latency between the kernels of an iteration, no
// host round trip to learn that every line has finished, and the loop ends as soon as it has (the reference
// keeps stepping finished lines and ignores them, seq2seq.py:1255-1256, so stopping early changes nothing).
// Kernel arguments are frozen into a graph, so one is valid for exactly one (workspace, geometry, parameters)
// combination; a handle keeps the last few (a page-by-page caller repeats few geometries, the bench one).
#pragma once
#include "model.h"
#include <vector>

namespace cor {

struct LoopGraph {
    std::vector<uint64_t> key;
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
    int body_kernels = 0;          // kernel nodes of one body
    std::vector<char> ctl;         // loop controls (conditional handles, thresholds) the graph's first node reads from memory
    uint64_t used = 0;
};

// Entry for `key` (created empty -- exec == nullptr -- if absent, evicting the least recently used one).
int loop_graph_lookup(cor_model* m, const std::vector<uint64_t>& key, LoopGraph** out);
// Stream to capture bodies on (the caller's stream may be the legacy default stream, which cannot capture).
// (`which` = 1: a second one, for the branches of a SWITCH node added while the first is capturing.)
cudaStream_t loop_capture_stream(cor_model* m, int which = 0);
void loop_graphs_release(cor_model* m);
// False while per-launch profiling is on (it brackets launches with events from the host) or COR_NO_GRAPH is set.
bool loop_graphs_enabled();

// WHILE node behind `first` (a kernel node that sets the initial condition): returns the body graph to capture into.
int loop_graph_begin(LoopGraph* g, const cudaKernelNodeParams& first, cudaGraphConditionalHandle* cond, cudaGraph_t* body);

// While `cs` is capturing: a SWITCH node on the selector `sw` (a handle of the top-level graph) with `n` branches behind
// everything captured so far; what is captured on `cs` afterwards depends on it.  Returns the branch graphs (to capture
// into with another stream).
// `n` == 0: an IF node (one branch, taken when the selector is non-zero) instead of a SWITCH.
int loop_graph_switch(cudaStream_t cs, int n, cudaGraphConditionalHandle sw, cudaGraph_t* branches);

}  // namespace cor
