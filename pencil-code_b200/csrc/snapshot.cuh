// snapshot.cuh -- var.dat / VAR<n> written from the device without stalling the time loop.
//
// The reference's wsnap (snapshot.f90:286-302) copies all of f to the host synchronously
// (copy_farray_from_GPU) and then writes it with output_snap (io_dist.f90:110-167).  Here the time loop
// only pays a device-to-device copy of f into a staging array (2.4 ms at 512^3 x 7); a worker thread pulls
// the staging array through two page-locked chunk buffers on its own stream (copy of chunk i+1 overlaps
// the write of chunk i, and both overlap the next time steps) and writes the reference's record layout:
//   record 1: f(mx,my,mz,mvar) incl. ghost zones     record 2: t, x(mx), y(my), z(mz), dx, dy, dz
// as Fortran sequential unformatted records with 4-byte markers, split gfortran-style above 2147483639
// bytes (negative leading marker = continued, negative trailing marker = has a predecessor).
#pragma once
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <string>
#include <thread>
#include <vector>

struct SnapWriter {
  double *dev = nullptr; size_t nelem = 0;                 // staging copy of f
  char *pin[2] = {nullptr, nullptr}; size_t chunk = 0;     // page-locked chunk buffers
  cudaStream_t st = nullptr;
  cudaEvent_t ev_copy = nullptr, ev_chunk[2] = {nullptr, nullptr};
  std::thread th; bool busy = false;
  int rc = 0; std::string err;
  int device = 0;

  // a caller that exits with a snapshot in flight (no wait, no finalize): finish the file instead of std::terminate
  ~SnapWriter() { if (th.joinable()) th.join(); }

  int fail_(const std::string &m) { rc = -4; err = m; return rc; }

  // worker: record 1 from the staging array, record 2 from `tail`
  void run(std::string path, std::vector<double> tail, long long rec_max)
  {
    rc = 0; err.clear();
    if (cudaSetDevice(device) != cudaSuccess) { fail_("snapshot: cudaSetDevice failed"); return; }
    FILE *fh = fopen(path.c_str(), "wb");
    if (!fh) { rc = -3; err = "snapshot: cannot open " + path; return; }
    if (cudaStreamWaitEvent(st, ev_copy, 0) != cudaSuccess) { fail_("snapshot: cudaStreamWaitEvent failed"); fclose(fh); return; }
    const char *src = reinterpret_cast<const char *>(dev);
    const long long total = (long long)nelem * 8;
    // pieces never cross a subrecord boundary; `lead`/`trail` markers are written around them
    struct Piece { long long off, len; int32_t lead, trail; bool has_lead, has_trail; };
    std::vector<Piece> pieces;
    {
      long long off = 0; bool first = true;
      while (true) {
        const long long k = std::min(total - off, rec_max);
        const bool more = off + k < total;
        long long o = off;
        while (o < off + k) {
          const long long len = std::min<long long>((long long)chunk, off + k - o);
          Piece p{o, len, 0, 0, o == off, o + len == off + k};
          p.lead = (int32_t)(more ? -k : k); p.trail = (int32_t)(first ? k : -k);
          pieces.push_back(p);
          o += len;
        }
        off += k; first = false;
        if (!more) break;
      }
    }
    auto issue = [&](size_t i) -> bool {
      const int b = (int)(i & 1);
      if (cudaMemcpyAsync(pin[b], src + pieces[i].off, (size_t)pieces[i].len, cudaMemcpyDeviceToHost, st) != cudaSuccess) return false;
      return cudaEventRecord(ev_chunk[b], st) == cudaSuccess;
    };
    bool ok = pieces.empty() || issue(0);
    for (size_t i = 0; ok && i < pieces.size(); i++) {
      const int b = (int)(i & 1);
      ok = cudaEventSynchronize(ev_chunk[b]) == cudaSuccess;
      if (ok && i + 1 < pieces.size()) ok = issue(i + 1);          // next copy runs while this chunk goes to disk
      if (!ok) break;
      const Piece &p = pieces[i];
      if (p.has_lead) ok = fwrite(&p.lead, 4, 1, fh) == 1;
      if (ok) ok = fwrite(pin[b], 1, (size_t)p.len, fh) == (size_t)p.len;
      if (ok && p.has_trail) ok = fwrite(&p.trail, 4, 1, fh) == 1;
    }
    if (ok) {     // record 2 (always far below the subrecord limit)
      const int32_t n = (int32_t)(tail.size() * 8);
      ok = fwrite(&n, 4, 1, fh) == 1 && fwrite(tail.data(), 8, tail.size(), fh) == tail.size() && fwrite(&n, 4, 1, fh) == 1;
    }
    if (fclose(fh) != 0) ok = false;
    if (!ok) { cudaGetLastError(); fail_("snapshot: copy or write failed for " + path); }
  }

  int wait()
  {
    if (busy) { th.join(); busy = false; }
    return rc;
  }
  void release()
  {
    wait();
    if (dev) cudaFree(dev); dev = nullptr; nelem = 0;
    for (int b = 0; b < 2; b++) { if (pin[b]) cudaFreeHost(pin[b]); pin[b] = nullptr; }
    chunk = 0;
  }
};
