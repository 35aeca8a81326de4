// AUC-family metric calculators (host side).
// Reference: BasicAucCalculator (fleet/metrics.h:46-196, metrics.cc:38-409: add_data
// variants, compute, calculate_bucket_error; WuAUC :501-587; NaN/Inf :452-475).
// The bucket tables can also be filled on the device (csrc/cuda/dense_ops.cu
// k_auc_bucket) and merged in with pbh_auc_merge_tables.
#include <algorithm>
#include <cmath>
#include <mutex>
#include <vector>

#include "util.h"

namespace pbh {

class AucCalculator {
 public:
  explicit AucCalculator(int table_size) { init(table_size); }
  void init(int table_size) {
    std::lock_guard<std::mutex> g(mu_);
    n_ = table_size;
    table_[0].assign(n_, 0.0);
    table_[1].assign(n_, 0.0);
    reset_locked();
  }
  void reset() { std::lock_guard<std::mutex> g(mu_); reset_locked(); }

  // label in {0,1} (or a float in [0,1] when float_label); mask (nullable) selects rows
  // float_label 2 = continuous values: error / mean sums only, nothing clamped or bucketed
  // (add_unlock_data_with_continue_label, fleet/metrics.cc:98-105)
  int add_data(const float* pred, const float* label, const float* mask, int64_t n, int float_label) {
    std::lock_guard<std::mutex> g(mu_);
    for (int64_t i = 0; i < n; ++i) {
      if (mask && mask[i] == 0.f) continue;
      double p = pred[i];
      if (float_label == 2) {
        double y = label[i];
        abserr_ += std::fabs(p - y);
        sqrerr_ += (p - y) * (p - y);
        pred_sum_ += p;
        label_sum_ += y;
        total_ += 1;
        continue;
      }
      if (!(p >= 0.0 && p <= 1.0)) {   // also catches NaN
        if (std::isnan(p)) nan_cnt_ += 1; else if (std::isinf(p)) inf_cnt_ += 1;
        p = std::min(1.0, std::max(0.0, std::isnan(p) ? 0.0 : p));
      }
      int pos = std::min((int)(p * n_), n_ - 1);
      double y = label[i];
      if (float_label) {
        table_[0][pos] += 1.0 - y;
        table_[1][pos] += y;
      } else {
        table_[y > 0.5 ? 1 : 0][pos] += 1.0;
        y = y > 0.5 ? 1.0 : 0.0;
      }
      abserr_ += std::fabs(p - y);
      sqrerr_ += (p - y) * (p - y);
      pred_sum_ += p;
      label_sum_ += y;
      total_ += 1;
    }
    return 0;
  }

  // per-user records for WuAUC
  void add_uid_data(const float* pred, const float* label, const uint64_t* uid, int64_t n) {
    std::lock_guard<std::mutex> g(mu_);
    for (int64_t i = 0; i < n; ++i) recs_.push_back({uid[i], label[i] > 0.5f ? 1 : 0, pred[i]});
  }

  void merge_tables(const double* tables /* [2, n] */, const double* sums /* [5] */) {
    std::lock_guard<std::mutex> g(mu_);
    for (int i = 0; i < n_; ++i) { table_[0][i] += tables[i]; table_[1][i] += tables[n_ + i]; }
    abserr_ += sums[0]; sqrerr_ += sums[1]; pred_sum_ += sums[2]; label_sum_ += sums[3]; total_ += sums[4];
  }
  // raw state, for cross-rank reduction by the caller: tables [2, n] then sums [5]
  void export_state(double* tables, double* sums) {
    std::lock_guard<std::mutex> g(mu_);
    std::copy(table_[0].begin(), table_[0].end(), tables);
    std::copy(table_[1].begin(), table_[1].end(), tables + n_);
    sums[0] = abserr_; sums[1] = sqrerr_; sums[2] = pred_sum_; sums[3] = label_sum_; sums[4] = total_;
  }
  void import_state(const double* tables, const double* sums) {
    std::lock_guard<std::mutex> g(mu_);
    std::copy(tables, tables + n_, table_[0].begin());
    std::copy(tables + n_, tables + 2 * n_, table_[1].begin());
    abserr_ = sums[0]; sqrerr_ = sums[1]; pred_sum_ = sums[2]; label_sum_ = sums[3]; total_ = sums[4];
  }

  // out: [auc, bucket_error, mae, rmse, actual_ctr, predicted_ctr, copc, size]  (box_wrapper.cc:1027-1046)
  void compute(double* out) {
    std::lock_guard<std::mutex> g(mu_);
    double area = 0, fp = 0, tp = 0;
    for (int i = n_ - 1; i >= 0; --i) {
      double nfp = fp + table_[0][i], ntp = tp + table_[1][i];
      area += (nfp - fp) * (tp + ntp) / 2;
      fp = nfp; tp = ntp;
    }
    double auc = (fp < 1e-3 || tp < 1e-3) ? -0.5 : area / (fp * tp);
    double size = fp + tp;
    double mae = size > 0 ? abserr_ / size : 0, rmse = size > 0 ? std::sqrt(sqrerr_ / size) : 0;
    double pctr = size > 0 ? pred_sum_ / size : 0, actr = size > 0 ? tp / size : 0;
    out[0] = auc; out[1] = bucket_error(); out[2] = mae; out[3] = rmse; out[4] = actr; out[5] = pctr;
    out[6] = pctr > 0 ? actr / pctr : 0; out[7] = size;
  }

  // out: [uauc, wuauc, user_cnt, size]
  void compute_wuauc(double* out) {
    std::lock_guard<std::mutex> g(mu_);
    std::sort(recs_.begin(), recs_.end(), [](const Rec& a, const Rec& b) {
      if (a.uid != b.uid) return a.uid > b.uid;
      if (a.pred != b.pred) return a.pred > b.pred;
      return a.label < b.label;
    });
    double users = 0, size = 0, uauc = 0, wuauc = 0;
    size_t i = 0;
    while (i < recs_.size()) {
      size_t j = i;
      while (j < recs_.size() && recs_[j].uid == recs_[i].uid) ++j;
      double tp = 0, fp = 0, area = 0;
      size_t k = i;
      while (k < j) {
        double ntp = tp, nfp = fp;
        size_t e = k;
        do { if (recs_[e].label == 1) ntp += 1; else nfp += 1; ++e; } while (e < j && recs_[e].pred == recs_[k].pred);
        area += (nfp - fp) * (tp + ntp) / 2.0;
        tp = ntp; fp = nfp;
        k = e;
      }
      if (tp > 0 && fp > 0) {
        double a = area / (fp * tp + 1e-9), ins = tp + fp;
        users += 1; size += ins; uauc += a; wuauc += a * ins;
      }
      i = j;
    }
    out[0] = users > 0 ? uauc / users : 0;
    out[1] = size > 0 ? wuauc / size : 0;
    out[2] = users; out[3] = size;
  }

  double nan_count() const { return nan_cnt_; }
  double inf_count() const { return inf_cnt_; }
  int table_size() const { return n_; }

 private:
  struct Rec { uint64_t uid; int label; float pred; };
  void reset_locked() {
    std::fill(table_[0].begin(), table_[0].end(), 0.0);
    std::fill(table_[1].begin(), table_[1].end(), 0.0);
    abserr_ = sqrerr_ = pred_sum_ = label_sum_ = total_ = 0;
    nan_cnt_ = inf_cnt_ = 0;
    recs_.clear();
  }
  double bucket_error() const {
    const double kRelErr = 0.05, kMaxSpan = 0.01;
    double last_ctr = -1, imp = 0, ctr_sum = 0, clk = 0, err_sum = 0, err_cnt = 0;
    for (int i = 0; i < n_; ++i) {
      double click = table_[1][i], show = table_[0][i] + table_[1][i];
      double ctr = (double)i / n_;
      if (std::fabs(ctr - last_ctr) > kMaxSpan) { last_ctr = ctr; imp = 0; ctr_sum = 0; clk = 0; }
      imp += show; ctr_sum += ctr * show; clk += click;
      double adj = ctr_sum / imp;
      double rel = std::sqrt((1 - adj) / (adj * imp));
      if (rel < kRelErr) {
        double actual = clk / imp;
        err_sum += std::fabs(actual / adj - 1) * imp;
        err_cnt += imp;
        last_ctr = -1;
      }
    }
    return err_cnt > 0 ? err_sum / err_cnt : 0.0;
  }
  std::mutex mu_;
  int n_ = 0;
  std::vector<double> table_[2];
  double abserr_ = 0, sqrerr_ = 0, pred_sum_ = 0, label_sum_ = 0, total_ = 0, nan_cnt_ = 0, inf_cnt_ = 0;
  std::vector<Rec> recs_;
};

}  // namespace pbh

using pbh::AucCalculator;
extern "C" {
int pbh_auc_create(void** out, int table_size) { *out = new AucCalculator(table_size); return 0; }
int pbh_auc_destroy(void* h) { delete (AucCalculator*)h; return 0; }
int pbh_auc_reset(void* h) { ((AucCalculator*)h)->reset(); return 0; }
int pbh_auc_add(void* h, const float* pred, const float* label, const float* mask, int64_t n, int float_label) {
  return ((AucCalculator*)h)->add_data(pred, label, mask, n, float_label);
}
int pbh_auc_add_uid(void* h, const float* pred, const float* label, const uint64_t* uid, int64_t n) {
  ((AucCalculator*)h)->add_uid_data(pred, label, uid, n);
  return 0;
}
int pbh_auc_merge_tables(void* h, const double* tables, const double* sums) {
  ((AucCalculator*)h)->merge_tables(tables, sums);
  return 0;
}
int pbh_auc_export(void* h, double* tables, double* sums) { ((AucCalculator*)h)->export_state(tables, sums); return 0; }
int pbh_auc_import(void* h, const double* tables, const double* sums) {
  ((AucCalculator*)h)->import_state(tables, sums);
  return 0;
}
int pbh_auc_compute(void* h, double* out8) { ((AucCalculator*)h)->compute(out8); return 0; }
int pbh_auc_compute_wuauc(void* h, double* out4) { ((AucCalculator*)h)->compute_wuauc(out4); return 0; }
int pbh_auc_nan_inf(void* h, double* out2) {
  out2[0] = ((AucCalculator*)h)->nan_count();
  out2[1] = ((AucCalculator*)h)->inf_count();
  return 0;
}
}
