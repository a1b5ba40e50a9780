// Native block executor: runs a list of commands - kernel launches through the library's own entry points, event
// records, stream waits, async copies - in ONE call from Python.
//
// A block iteration is ~45 launches on three streams.  Issued one by one through ctypes they cost the host ~1 ms per
// block, which is more than the device needs for the block's whole dependency chain; issued from here they cost a few
// microseconds each.  The reference's counterpart is the Go peer's scheduler loop (prepareForNextIteration and the
// timer goroutines, DistSys/main.go:1059-1187, 1494-1589): native there, native here.  What runs when stays readable
// in one place on the Python side (protocol/schedule.py builds the lists once and patches the few per-block words);
// this file only dispatches.  Opcodes >= 0 are generated from the header (tools/gen_dispatch.py -> dispatch_gen.inc).
#include "common.cuh"
#include "../../include/biscotti_b200.h"
#include <string.h>

namespace {

inline float u2f(uint64_t v) {
  uint32_t w = (uint32_t)v;
  float f;
  memcpy(&f, &w, 4);
  return f;
}
inline double u2d(uint64_t v) {
  double d;
  memcpy(&d, &v, 8);
  return d;
}

enum SpecialOp {
  OP_NOP = -1,
  OP_EVENT_RECORD = -2,   // a0 = event
  OP_STREAM_WAIT = -3,    // a0 = event
  OP_MEMCPY_D2H = -4,     // a0 = dst (pinned host), a1 = src (device), a2 = bytes
  OP_MEMCPY_H2D = -5,     // a0 = dst (device), a1 = src (pinned host), a2 = bytes
  OP_MEMSET = -6,         // a0 = ptr, a1 = byte value, a2 = bytes
  OP_MEMCPY_D2D = -7,     // a0 = dst, a1 = src, a2 = bytes
};

int dispatch(int op, const uint64_t* a, void* st) {
  switch (op) {
#include "dispatch_gen.inc"
    default:
      return -1000;
  }
}

}  // namespace

extern "C" {

struct BscCmd {
  int32_t op;
  int32_t stream;       // index into the `streams` array given to bsc_exec
  uint64_t a[24];
};

// Runs cmds[0..n) in order.  Returns 0, or (index + 1) * 100000 + |rc| of the first failing command.
int bsc_exec(const void* cmds_, int n, void* const* streams, int nstreams) {
  const BscCmd* cmds = (const BscCmd*)cmds_;
  for (int i = 0; i < n; i++) {
    const BscCmd& c = cmds[i];
    if (c.op == OP_NOP) continue;
    if (c.stream < 0 || c.stream >= nstreams) return (i + 1) * 100000 + 999;
    cudaStream_t st = (cudaStream_t)streams[c.stream];
    int rc = 0;
    switch (c.op) {
      case OP_EVENT_RECORD:
        rc = (int)cudaEventRecord((cudaEvent_t)c.a[0], st);
        break;
      case OP_STREAM_WAIT:
        rc = (int)cudaStreamWaitEvent(st, (cudaEvent_t)c.a[0], 0);
        break;
      case OP_MEMCPY_D2H:
        rc = (int)cudaMemcpyAsync((void*)c.a[0], (const void*)c.a[1], (size_t)c.a[2], cudaMemcpyDeviceToHost, st);
        break;
      case OP_MEMCPY_H2D:
        rc = (int)cudaMemcpyAsync((void*)c.a[0], (const void*)c.a[1], (size_t)c.a[2], cudaMemcpyHostToDevice, st);
        break;
      case OP_MEMCPY_D2D:
        rc = (int)cudaMemcpyAsync((void*)c.a[0], (const void*)c.a[1], (size_t)c.a[2], cudaMemcpyDeviceToDevice, st);
        break;
      case OP_MEMSET:
        rc = (int)cudaMemsetAsync((void*)c.a[0], (int)c.a[1], (size_t)c.a[2], st);
        break;
      default:
        rc = dispatch(c.op, c.a, (void*)st);
    }
    if (rc != 0) {
      fprintf(stderr, "[biscotti_b200] bsc_exec: command %d (op %d) failed with %d\n", i, c.op, rc);
      return (i + 1) * 100000 + (rc < 0 ? -rc : rc);
    }
  }
  return 0;
}
int bsc_cmd_sizeof() { return (int)sizeof(BscCmd); }

// Events the command lists record and wait on (timing disabled: cheapest form).
void* bsc_event_create() {
  cudaEvent_t e = nullptr;
  if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) return nullptr;
  return (void*)e;
}
void bsc_event_destroy(void* e) { cudaEventDestroy((cudaEvent_t)e); }
int bsc_event_record(void* e, void* stream) { return (int)cudaEventRecord((cudaEvent_t)e, (cudaStream_t)stream); }
int bsc_event_synchronize(void* e) { return (int)cudaEventSynchronize((cudaEvent_t)e); }
int bsc_event_query(void* e) { return cudaEventQuery((cudaEvent_t)e) == cudaSuccess ? 1 : 0; }
int bsc_stream_wait_event(void* stream, void* e) { return (int)cudaStreamWaitEvent((cudaStream_t)stream, (cudaEvent_t)e, 0); }

}  // extern "C"
