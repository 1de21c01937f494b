// dk_api_multi.cu — one model replicated on several GPUs of THIS process (include/dk_b200.h, dk_multi_*).
//
// R cannot fork CUDA contexts (the reference's only parallelism, mclapply over restarts at
// /root/reference/R/max_crit.R:125, does not survive on a GPU), so the split over devices lives inside the library:
// candidates (or qEI batches) are cut into contiguous ranges, one per device, sizes differing by at most one; one host
// thread per device runs the chunked host <-> device pipeline on its range of the caller's arrays (no copy of the
// whole matrix, no device-to-device traffic); the only merge is the host-side selection of the k best
// (value, global index) pairs, ties going to the lower index, i.e. which.max's "first maximum".
#include <string>
#include <thread>

#include "dk_host.h"

using namespace dk;

struct dk_multi {
  std::vector<dk_model*> models;
};

namespace {

void shard_range(int64_t N, int r, int world, int64_t* lo, int64_t* hi) {
  const int64_t base = N / world, rem = N % world;
  *lo = r * base + std::min<int64_t>(r, rem);
  *hi = *lo + base + (r < rem ? 1 : 0);
}

// Run fn(replica index) on one host thread per device; the first failure's status and message win.
template <class Fn>
int on_every_device(dk_multi* mm, Fn fn) {
  const int G = (int)mm->models.size();
  std::vector<int> rc(G, DK_OK);
  std::vector<std::string> msg(G);
  auto body = [&](int g) {
    rc[g] = fn(g);
    if (rc[g]) msg[g] = dk_last_error();   // the message is thread-local: carry it to the caller's thread
  };
  if (G == 1) {
    body(0);
  } else {
    std::vector<std::thread> th;
    th.reserve(G);
    for (int g = 0; g < G; g++) th.emplace_back(body, g);
    for (auto& t : th) t.join();
  }
  for (int g = 0; g < G; g++)
    if (rc[g]) return fail(rc[g], "device %d: %s", mm->models[g]->device, msg[g].c_str());
  return DK_OK;
}

int check_multi(const dk_multi* mm, const void* x, int64_t N) {
  if (!mm || mm->models.empty()) return fail(DK_ERR_ARG, "multi-GPU handle is NULL");
  return check_call(mm->models[0], x, N);
}

}  // namespace

extern "C" {

int dk_model_clone(const dk_model* src, int device, dk_model** out);

int dk_multi_create(const dk_model_desc* desc, const int* devices, int ndev, dk_multi** out) {
  if (!desc || !out) return fail(DK_ERR_ARG, "dk_multi_create: NULL argument");
  *out = nullptr;
  int visible = 0;
  if (cudaGetDeviceCount(&visible) != cudaSuccess || visible == 0) return fail(DK_ERR_UNSUPPORTED, "no CUDA device visible");
  std::vector<int> devs;
  if (!devices) {
    if (ndev <= 0 || ndev > visible) ndev = visible;   // NULL list: the first ndev devices (all of them for ndev <= 0)
    for (int g = 0; g < ndev; g++) devs.push_back(g);
  } else {
    if (ndev < 1) return fail(DK_ERR_ARG, "empty device list");
    for (int g = 0; g < ndev; g++) {
      if (devices[g] < 0 || devices[g] >= visible) return fail(DK_ERR_ARG, "device %d out of range (%d visible)", devices[g], visible);
      devs.push_back(devices[g]);
    }
  }
  dk_multi* mm = new (std::nothrow) dk_multi();
  if (!mm) return fail(DK_ERR_NOMEM, "host allocation failed");
  mm->models.assign(devs.size(), nullptr);
  dk_model_desc d0 = *desc;
  d0.device = devs[0];
  int rc = dk_model_create(&d0, &mm->models[0]);   // tiles the factor once, on the host
  if (!rc) {
    // replicas: device-to-device copies of the tiled model, in parallel
    std::vector<int> rcs(devs.size(), DK_OK);
    std::vector<std::string> msgs(devs.size());
    std::vector<std::thread> th;
    for (size_t g = 1; g < devs.size(); g++)
      th.emplace_back([&, g] {
        rcs[g] = dk_model_clone(mm->models[0], devs[g], &mm->models[g]);
        if (rcs[g]) msgs[g] = dk_last_error();
      });
    for (auto& t : th) t.join();
    for (size_t g = 1; g < devs.size() && !rc; g++)
      if (rcs[g]) rc = fail(rcs[g], "device %d: %s", devs[g], msgs[g].c_str());
  }
  if (rc) {
    const std::string keep = dk_last_error();
    for (dk_model* m : mm->models) dk_model_destroy(m);
    delete mm;
    return fail(rc, "%s", keep.c_str());
  }
  *out = mm;
  return DK_OK;
}

void dk_multi_destroy(dk_multi* mm) {
  if (!mm) return;
  for (dk_model* m : mm->models) dk_model_destroy(m);
  delete mm;
}

int dk_multi_count(const dk_multi* mm) { return mm ? (int)mm->models.size() : 0; }
dk_model* dk_multi_model(dk_multi* mm, int i) {
  return (mm && i >= 0 && i < (int)mm->models.size()) ? mm->models[i] : nullptr;
}

int dk_multi_append(dk_multi* mm, const double* x_new, double y_new, double new_noise_var) {
  if (!mm || mm->models.empty()) return fail(DK_ERR_ARG, "multi-GPU handle is NULL");
  if (!x_new) return fail(DK_ERR_ARG, "x_new is NULL");
  // every replica grows its own copy (the n^2 host substitution is repeated per device: a few ms each, in parallel)
  return on_every_device(mm, [&](int g) { return dk_model_append(mm->models[g], x_new, y_new, new_noise_var); });
}

int dk_multi_predict(dk_multi* mm, const double* x, int64_t N, int type, double* out_mean, double* out_sd) {
  int rc = check_multi(mm, x, N);
  if (rc) return rc;
  if (!out_mean) return fail(DK_ERR_ARG, "out_mean is NULL");
  const int G = (int)mm->models.size();
  CritParams P{MODE_PREDICT, type, 1, 0, 0.0, 0.0};
  return on_every_device(mm, [&](int g) {
    int64_t lo, hi;
    shard_range(N, g, G, &lo, &hi);
    return host_predict(mm->models[g], P, x, N, lo, hi - lo, nullptr, out_mean, out_sd);
  });
}

int dk_multi_ei(dk_multi* mm, const double* x, int64_t N, double plugin, int type, int minimization, int proxy,
                double* out_ei, double* out_mean, double* out_sd) {
  int rc = check_multi(mm, x, N);
  if (rc) return rc;
  if (!out_ei) return fail(DK_ERR_ARG, "out_ei is NULL");
  const int G = (int)mm->models.size();
  CritParams P{MODE_EI, type, minimization, proxy, plugin, 0.0};
  return on_every_device(mm, [&](int g) {
    int64_t lo, hi;
    shard_range(N, g, G, &lo, &hi);
    return host_predict(mm->models[g], P, x, N, lo, hi - lo, out_ei, out_mean, out_sd);
  });
}

int dk_multi_ei_grad(dk_multi* mm, const double* x, int64_t N, double plugin, int type, int minimization, int proxy,
                     double* out_grad, double* out_ei) {
  int rc = check_multi(mm, x, N);
  if (rc) return rc;
  if (!out_grad) return fail(DK_ERR_ARG, "out_grad is NULL");
  const int G = (int)mm->models.size();
  CritParams P{MODE_EI, type, minimization, proxy, plugin, 0.0};
  return on_every_device(mm, [&](int g) {
    int64_t lo, hi;
    shard_range(N, g, G, &lo, &hi);
    return host_grad(mm->models[g], P, x, N, lo, hi - lo, out_grad, out_ei);
  });
}

int dk_multi_quantile(dk_multi* mm, const double* x, int64_t N, double qnorm_beta, int type, double* out_q,
                      double* out_grad) {
  int rc = check_multi(mm, x, N);
  if (rc) return rc;
  if (!out_q) return fail(DK_ERR_ARG, "out_q is NULL");
  const int G = (int)mm->models.size();
  CritParams P{MODE_QUANTILE, type, 1, 0, 0.0, qnorm_beta};
  return on_every_device(mm, [&](int g) {
    int64_t lo, hi;
    shard_range(N, g, G, &lo, &hi);
    if (out_grad) return host_grad(mm->models[g], P, x, N, lo, hi - lo, out_grad, out_q);
    return host_predict(mm->models[g], P, x, N, lo, hi - lo, out_q, nullptr, nullptr);
  });
}

int dk_multi_akg(dk_multi* mm, const double* x, int64_t N, double new_noise_var, int type, double* out) {
  int rc = check_multi(mm, x, N);
  if (rc) return rc;
  if (!out) return fail(DK_ERR_ARG, "out is NULL");
  const int G = (int)mm->models.size();
  return on_every_device(mm, [&](int g) {
    int64_t lo, hi;
    shard_range(N, g, G, &lo, &hi);
    return host_akg(mm->models[g], x, N, lo, hi - lo, new_noise_var, type, out);
  });
}

int dk_multi_qei(dk_multi* mm, const double* x, int64_t B, int q, double plugin, int type, int minimization,
                 double* out, double* out_grad) {
  int rc = check_multi(mm, x, B);
  if (rc) return rc;
  if (q < 1 || q > DK_MAX_Q_LATTICE) return fail(DK_ERR_ARG, "qEI batch size q=%d outside 1..%d", q, DK_MAX_Q_LATTICE);
  if (out_grad && q < 2) return fail(DK_ERR_ARG, "qEI.grad needs q >= 2 (use dk_multi_ei_grad for q = 1)");
  if (!out && !out_grad) return fail(DK_ERR_ARG, "both outputs are NULL");
  const int G = (int)mm->models.size();
  return on_every_device(mm, [&](int g) {
    int64_t lo, hi;
    shard_range(B, g, G, &lo, &hi);
    return host_qei(mm->models[g], x, B * q, lo, hi - lo, q, plugin, type, minimization, out, out_grad);
  });
}

namespace {
int multi_screen(dk_multi* mm, const ScreenSpec& S, const double* x, const dk_uniform_spec* gen, int64_t N, int k,
                 int64_t* out_idx, double* out_val) {
  const int G = (int)mm->models.size();
  std::vector<int64_t> idx((size_t)G * k, -1);
  std::vector<double> val((size_t)G * k, 0.0);
  int rc = on_every_device(mm, [&](int g) {
    int64_t lo, hi;
    shard_range(N, g, G, &lo, &hi);
    int r = host_screen(mm->models[g], S, x, 0, gen, N, lo, hi - lo, k, &idx[(size_t)g * k], &val[(size_t)g * k], nullptr);
    if (!r)
      for (int i = 0; i < k; i++)
        if (idx[(size_t)g * k + i] >= 0) idx[(size_t)g * k + i] += lo;   // shard-local -> global row
    return r;
  });
  if (rc) return rc;
  // merge: best first (EI, AKG: larger; quantile: smaller), then index ascending (which.max keeps the first maximum)
  const double sign = S.criterion == DK_CRIT_QUANTILE ? -1.0 : 1.0;
  std::vector<int> order;
  for (int i = 0; i < G * k; i++)
    if (idx[i] >= 0) order.push_back(i);
  std::sort(order.begin(), order.end(), [&](int a, int b) {
    if (val[a] != val[b]) return sign * val[a] > sign * val[b];
    return idx[a] < idx[b];
  });
  for (int i = 0; i < k; i++) {
    const bool have = i < (int)order.size();
    out_idx[i] = have ? idx[order[i]] : -1;
    out_val[i] = have ? val[order[i]] : 0.0;
  }
  return DK_OK;
}
}  // namespace

int dk_multi_screen_ei(dk_multi* mm, const double* x, int64_t N, double plugin, int type, int minimization, int proxy,
                       int k, int64_t* out_idx, double* out_val) {
  int rc = check_multi(mm, x, N);
  if (rc) return rc;
  if (k < 1 || k > 64) return fail(DK_ERR_ARG, "k=%d outside 1..64", k);
  if (!out_idx || !out_val) return fail(DK_ERR_ARG, "output pointer is NULL");
  ScreenSpec S{DK_CRIT_EI, CritParams{MODE_EI, type, minimization, proxy, plugin, 0.0}, 0.0};
  return multi_screen(mm, S, x, nullptr, N, k, out_idx, out_val);
}

int dk_multi_screen(dk_multi* mm, const double* x, const dk_uniform_spec* gen, int64_t N, const dk_crit_spec* crit, int k,
                    int64_t* out_idx, double* out_val, double* out_x) {
  if (!mm || mm->models.empty()) return fail(DK_ERR_ARG, "multi-GPU handle is NULL");
  if (!x && !gen) return fail(DK_ERR_ARG, "neither candidates nor a generator were given");
  if (N < 0) return fail(DK_ERR_ARG, "N = %lld is negative", (long long)N);
  if (k < 1 || k > 64) return fail(DK_ERR_ARG, "k=%d outside 1..64", k);
  if (!out_idx || !out_val) return fail(DK_ERR_ARG, "output pointer is NULL");
  if (out_x && x) return fail(DK_ERR_ARG, "out_x is for generated candidates (the caller already holds x)");
  ScreenSpec S;
  int rc = make_screen_spec(mm->models[0], crit, &S);
  if (rc) return rc;
  if ((rc = multi_screen(mm, S, x, x ? nullptr : gen, N, k, out_idx, out_val))) return rc;
  if (out_x) generated_rows(gen, mm->models[0]->d, out_idx, k, out_x);
  return DK_OK;
}

}  // extern "C"
