// Keyed CUDA-graph cache for host-side launch sequences (csrc/backward.cu; forward.cu keeps its own, older copy of the
// same policy): the caller serialises every byte that reaches a launch into a key; the first sighting of a key runs the
// plain launches (which also sets kernel attributes outside any capture), the second captures + instantiates, later ones
// replay.  Per host thread; least-recently-used eviction.
#pragma once
#include "common.cuh"
#include <string.h>
#include <vector>

namespace skg {

struct KeyedGraph {
  std::vector<uint8_t> key;
  int seen = 0;
  cudaGraphExec_t exec = nullptr;
  long long launches = 0;
  unsigned long long stamp = 0;
};

struct GraphCache {
  static constexpr int kEntries = 16;
  KeyedGraph e[kEntries];
  unsigned long long clock = 0;
  long long stats[3] = {0, 0, 0};   // replays, captures, plain
  // the legacy default stream cannot be captured: work arriving on it runs on this library stream, forked from / joined
  // into the caller's stream with two events
  int run_dev = -1;
  cudaStream_t run_stream = nullptr;
  cudaEvent_t fork_ev = nullptr, join_ev = nullptr;
};

struct KeyBuilder {
  std::vector<uint8_t> bytes;
  template <class T> void add(const T& v) { const uint8_t* p = reinterpret_cast<const uint8_t*>(&v); bytes.insert(bytes.end(), p, p + sizeof(T)); }
  void add_raw(const void* p, size_t n) { const uint8_t* q = reinterpret_cast<const uint8_t*>(p); bytes.insert(bytes.end(), q, q + n); }
};

long long launch_count_now();   // = skg_launch_count()

// enqueue(stream, capturing) -> int (SKG_OK or an error); returns the enqueue's status
template <class Fn>
static int graph_cached_run(GraphCache& gc, const std::vector<uint8_t>& key, cudaStream_t st, Fn enqueue) {
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  if (st == nullptr || st == cudaStreamLegacy) {
    int dev = 0;
    SKG_CUDA(cudaGetDevice(&dev));
    if (gc.run_dev != dev) {
      if (gc.run_stream) { cudaStreamDestroy(gc.run_stream); cudaEventDestroy(gc.fork_ev); cudaEventDestroy(gc.join_ev); gc.run_stream = nullptr; }
      SKG_CUDA(cudaStreamCreateWithFlags(&gc.run_stream, cudaStreamNonBlocking));
      SKG_CUDA(cudaEventCreateWithFlags(&gc.fork_ev, cudaEventDisableTiming));
      SKG_CUDA(cudaEventCreateWithFlags(&gc.join_ev, cudaEventDisableTiming));
      gc.run_dev = dev;
    }
    SKG_CUDA(cudaEventRecord(gc.fork_ev, st));
    SKG_CUDA(cudaStreamWaitEvent(gc.run_stream, gc.fork_ev, 0));
    const int rc = graph_cached_run(gc, key, gc.run_stream, enqueue);
    SKG_CUDA(cudaEventRecord(gc.join_ev, gc.run_stream));
    SKG_CUDA(cudaStreamWaitEvent(st, gc.join_ev, 0));
    return rc;
  }
  if (cudaStreamIsCapturing(st, &cs) != cudaSuccess || cs != cudaStreamCaptureStatusNone)
    return enqueue(st, cs != cudaStreamCaptureStatusNone);
  KeyedGraph* hit = nullptr; KeyedGraph* victim = &gc.e[0];
  for (auto& g : gc.e) {
    if (g.seen > 0 && g.key == key) { hit = &g; break; }
    if (g.stamp < victim->stamp) victim = &g;
  }
  if (hit && hit->exec) {
    hit->stamp = ++gc.clock;
    SKG_CUDA(cudaGraphLaunch(hit->exec, st));
    add_launches(hit->launches);
    gc.stats[0]++;
    return SKG_OK;
  }
  if (!hit) {
    if (victim->exec) { cudaGraphExecDestroy(victim->exec); victim->exec = nullptr; }
    victim->key = key; victim->seen = 1; victim->stamp = ++gc.clock;
    gc.stats[2]++;
    return enqueue(st, false);
  }
  hit->stamp = ++gc.clock;
  const long long l0 = launch_count_now();
  if (cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal) != cudaSuccess) {
    cudaGetLastError();
    return enqueue(st, false);
  }
  const int rc = enqueue(st, true);
  cudaGraph_t graph = nullptr;
  const cudaError_t ce = cudaStreamEndCapture(st, &graph);
  const long long captured = launch_count_now() - l0;
  add_launches(-captured);
  if (rc != SKG_OK || ce != cudaSuccess || !graph) {
    cudaGetLastError();
    if (graph) cudaGraphDestroy(graph);
    hit->seen = 0;
    if (rc != SKG_OK) return rc;
    return enqueue(st, false);
  }
  cudaGraphExec_t exec = nullptr;
  const cudaError_t ie = cudaGraphInstantiate(&exec, graph, 0);
  cudaGraphDestroy(graph);
  if (ie != cudaSuccess || !exec) {
    cudaGetLastError();
    hit->seen = 0;
    return enqueue(st, false);
  }
  hit->exec = exec; hit->launches = captured;
  gc.stats[1]++;
  SKG_CUDA(cudaGraphLaunch(exec, st));
  add_launches(captured);
  return SKG_OK;
}

}  // namespace skg
