// transport_abi.cuh -- the shared-memory wire format of the reference's asynchronous MPC_Wrapper
// (/root/reference/scripts/MPC_Wrapper.py:47-52,227-264) served for MANY robots by one GPU.
// Included at the end of qrw_b200.cu.
//
// Per robot slot, exactly what MPC_Wrapper.compress_dataIn / convert_dataOut put in their multiprocessing Arrays:
//   dataIn  [1 + 12 (n+1) + 12 N_gait] doubles: k / k_mpc | xref in C order (12 rows of n+1) | fsteps in C order (N_gait rows of 12)
//   dataOut [24 n] doubles: the 24 x n result in F order (= this library's result layout)
//   newData / newResult: one signed char each ('b' Values)
// The server side (create_MPC_asynchronous, :170-225) becomes: clear newData, take the slot's dataIn, solve, write dataOut,
// set newResult -- for every slot that has new data, in one launch of the indexed MPC kernel.
#pragma once

namespace {
// wire -> ABI layout for the slots that are called in this round; one CTA per slot.
__global__ void __launch_bounds__(128) mpc_wire_ingest_kernel(const int B, const int n1, const int N_gait, const double* __restrict__ wire,
                                                             const int* __restrict__ k_arr, double* __restrict__ xref,
                                                             double* __restrict__ fsteps) {
  const int b = blockIdx.x;
  if (b >= B || k_arr[b] < 0) return;
  const int per = 1 + 12 * n1 + 12 * N_gait;
  const double* w = wire + (size_t)b * per + 1;
  double* X = xref + (size_t)b * 12 * n1;
  double* F = fsteps + (size_t)b * 12 * N_gait;
  for (int e = threadIdx.x; e < 12 * n1; e += blockDim.x) {  // C order [12][n1] -> column-major r + 12 c
    const int r = e / n1, c = e - r * n1;
    X[r + 12 * c] = w[e];
  }
  const double* wf = w + 12 * n1;
  for (int e = threadIdx.x; e < 12 * N_gait; e += blockDim.x) {  // C order [N_gait][12] -> column-major i + N_gait j
    const int i = e / 12, j = e - 12 * i;
    const double v = wf[e];
    F[i + N_gait * j] = (v != v) ? 0.0 : v;  // fsteps[np.isnan(fsteps)] = 0.0 (MPC_Wrapper.py:203,236)
  }
}
}  // namespace

struct qrw_mpc_server {
  qrw_mpc_t* mpc = nullptr;
  int B = 0, per_in = 0, per_out = 0;
  double* d_wire = nullptr;
  int* d_k = nullptr;
  double *h_wire = nullptr, *h_out = nullptr;  // pinned staging
  int* h_k = nullptr;
  unsigned char* seen = nullptr;  // per slot: has this robot been served before?  (its first request is a cold start whatever k it posts)
  long long served = 0, rounds = 0;
};

extern "C" {

int qrw_mpc_server_create(qrw_mpc_server_t** out, qrw_mpc_t* mpc) {
  if (!out || !mpc) return fail("qrw_mpc_server_create: null argument");
  if (mpc->variant != 2) return fail("qrw_mpc_server_create: needs the default MPC kernel (variant 2)");
  CU(cudaSetDevice(mpc->device));
  auto* s = new qrw_mpc_server();
  s->mpc = mpc, s->B = mpc->B;
  s->per_in = 1 + 12 * (NS + 1) + 12 * mpc->N_gait, s->per_out = 24 * NS;
  const size_t B = (size_t)s->B;
  CUH(cudaMalloc(&s->d_wire, B * s->per_in * sizeof(double)), qrw_mpc_server_destroy(s));
  CUH(cudaMalloc(&s->d_k, B * sizeof(int)), qrw_mpc_server_destroy(s));
  CUH(cudaMallocHost(&s->h_wire, B * s->per_in * sizeof(double)), qrw_mpc_server_destroy(s));
  CUH(cudaMallocHost(&s->h_out, B * s->per_out * sizeof(double)), qrw_mpc_server_destroy(s));
  CUH(cudaMallocHost(&s->h_k, B * sizeof(int)), qrw_mpc_server_destroy(s));
  s->seen = new unsigned char[B]();
  *out = s;
  return 0;
}
int qrw_mpc_server_destroy(qrw_mpc_server_t* s) {
  if (!s) return 0;
  cudaFree(s->d_wire), cudaFree(s->d_k), cudaFreeHost(s->h_wire), cudaFreeHost(s->h_out), cudaFreeHost(s->h_k);
  delete[] s->seen;
  delete s;
  return 0;
}
int qrw_mpc_server_slot_doubles(qrw_mpc_server_t* s, int* per_in, int* per_out) {
  if (!s) return fail("qrw_mpc_server_slot_doubles: null handle");
  if (per_in) *per_in = s->per_in;
  if (per_out) *per_out = s->per_out;
  return 0;
}

// One round of create_MPC_asynchronous' loop body (MPC_Wrapper.py:189-223) for every slot: slots whose newData flag is set
// are claimed (flag cleared FIRST, then dataIn read, as the reference does), solved in one launch and published (dataOut
// written, THEN newResult set).  data_in [B][per_in], data_out [B][per_out], flags [B] signed chars, all host pointers
// (normally inside one shared-memory block that the robot processes map too).  *n_served = number of slots solved.
int qrw_mpc_server_poll_host(qrw_mpc_server_t* s, const double* data_in, signed char* new_data, double* data_out,
                             signed char* new_result, int* n_served) {
  if (!s || !data_in || !new_data || !data_out || !new_result) return fail("qrw_mpc_server_poll_host: null argument");
  qrw_mpc_t* h = s->mpc;
  const int B = s->B;
  int n = 0;
  for (int b = 0; b < B; ++b) {
    s->h_k[b] = -1;
    if (__atomic_load_n(&new_data[b], __ATOMIC_ACQUIRE)) {
      // The reference's Arrays are lock-wrapped; here a client that posts AGAIN while its slot is being copied would leave a
      // torn xref / fsteps.  Clear the flag, copy, and copy again if the flag came back during the copy (the newer post wins).
      for (int tries = 0; tries < 8; ++tries) {
        __atomic_store_n(&new_data[b], (signed char)0, __ATOMIC_RELEASE);
        memcpy(s->h_wire + (size_t)b * s->per_in, data_in + (size_t)b * s->per_in, s->per_in * sizeof(double));
        if (!__atomic_load_n(&new_data[b], __ATOMIC_ACQUIRE)) break;
      }
      s->h_k[b] = (int)s->h_wire[(size_t)b * s->per_in];  // int(kf[0]), MPC_Wrapper.py:197
      // a slot's FIRST request is a cold start (create_matrices / osqp_setup) whatever index it carries: a robot attaching
      // mid-run, or a restarted server, would otherwise run the update path on an all-zero solver state (rho = 0 -> NaN)
      if (s->h_k[b] < 0 || !s->seen[b]) s->h_k[b] = 0;
      s->seen[b] = 1;
      ++n;
    }
  }
  ++s->rounds;
  if (n_served) *n_served = n;
  if (n == 0) return 0;
  CU(cudaSetDevice(h->device));
  cudaStream_t st = h->stream;
  // contiguous runs of claimed slots travel as one copy each
  for (int b = 0; b < B;) {
    if (s->h_k[b] < 0) { ++b; continue; }
    int e = b;
    while (e < B && s->h_k[e] >= 0) ++e;
    CU(cudaMemcpyAsync(s->d_wire + (size_t)b * s->per_in, s->h_wire + (size_t)b * s->per_in, (size_t)(e - b) * s->per_in * sizeof(double),
                       cudaMemcpyHostToDevice, st));
    b = e;
  }
  CU(cudaMemcpyAsync(s->d_k, s->h_k, B * sizeof(int), cudaMemcpyHostToDevice, st));
  mpc_wire_ingest_kernel<<<B, 128, 0, st>>>(B, NS + 1, h->N_gait, s->d_wire, s->d_k, h->d_xref, h->d_fsteps);
  ++g_launches;
  CU(cudaGetLastError());
  if (qrw_mpc_run_indexed_device(h, s->d_k, h->d_xref, h->d_fsteps, st)) return 1;
  for (int b = 0; b < B;) {
    if (s->h_k[b] < 0) { ++b; continue; }
    int e = b;
    while (e < B && s->h_k[e] >= 0) ++e;
    CU(cudaMemcpyAsync(s->h_out + (size_t)b * s->per_out, h->d_out + (size_t)b * s->per_out, (size_t)(e - b) * s->per_out * sizeof(double),
                       cudaMemcpyDeviceToHost, st));
    b = e;
  }
  CU(cudaStreamSynchronize(st));
  for (int b = 0; b < B; ++b)
    if (s->h_k[b] >= 0) {
      memcpy(data_out + (size_t)b * s->per_out, s->h_out + (size_t)b * s->per_out, s->per_out * sizeof(double));
      __atomic_store_n(&new_result[b], (signed char)1, __ATOMIC_RELEASE);
    }
  s->served += n;
  return 0;
}

// The infinite loop of create_MPC_asynchronous (`while running.value`), natively: polls until *running becomes 0 (or
// max_rounds > 0 rounds have been done), sleeping idle_us microseconds after a round that found nothing.
int qrw_mpc_server_loop_host(qrw_mpc_server_t* s, const double* data_in, signed char* new_data, double* data_out,
                             signed char* new_result, const signed char* running, int idle_us, long long max_rounds,
                             long long* total_served) {
  if (!s || !running) return fail("qrw_mpc_server_loop_host: null argument");
  long long rounds = 0;
  const long long served0 = s->served;
  while (__atomic_load_n(running, __ATOMIC_ACQUIRE)) {
    int n = 0;
    if (qrw_mpc_server_poll_host(s, data_in, new_data, data_out, new_result, &n)) return 1;
    if (n == 0 && idle_us > 0) {
      struct timespec ts = {0, (long)idle_us * 1000L};
      nanosleep(&ts, nullptr);
    }
    if (max_rounds > 0 && ++rounds >= max_rounds) break;
  }
  if (total_served) *total_served = s->served - served0;
  return 0;
}

}  // extern "C"
