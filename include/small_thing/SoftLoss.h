// small_thing/SoftLoss.h — loss descriptors of the host mirror.  Same type names as the reference's
// lib/small_thing/SoftLoss.h:16-176 for the two losses ark_vi_ba uses (viba/common/Constants.h:21-25):
// TrivialLoss and HuberLossWithCutoff(a, k).  The optimizer keeps a POINTER to the loss, as the
// reference does (Factor.h:141), so parameters changed after addFactor (VisualFactors.cpp:21,
// setSize) are the ones the device sees: they are read when the graph is handed to the C ABI.
// val() is provided for host-side bookkeeping; the device evaluates rho and rho' itself.
#pragma once

#include <cmath>
#include <limits>
#include <stdexcept>

namespace small_thing {

struct Loss {
  virtual ~Loss() = default;
  // (a, k) of the C ABI's vb_factor_store::loss_a/loss_k; a <= 0 or infinite means "identity"
  virtual double abiA() const = 0;
  virtual double abiK() const = 0;
};

struct TrivialLoss : Loss {
  explicit TrivialLoss(double /* a */ = 0) {}
  void setSize(double /* a */) {}
  static double val(double s) { return s; }
  double abiA() const override { return 0.0; }
  double abiK() const override { return 0.0; }
};

// rho(s) = s for s <= a^2 ; 2 a sqrt(s) - a^2 up to s = k^2 ; constant above (SoftLoss.h:131-163)
struct HuberLossWithCutoff : Loss {
  double a, k;
  explicit HuberLossWithCutoff(double a_, double k_) : a(a_), k(k_) { check(a_, k_); }
  void setSize(double a_, double k_) {
    check(a_, k_);
    a = a_;
    k = k_;
  }
  double val(double s) const {
    if (s <= a * a) return s;
    if (s <= k * k) return 2.0 * a * std::sqrt(s) - a * a;
    return 2.0 * a * k - a * a;
  }
  double abiA() const override { return a; }
  double abiK() const override { return k; }

 private:
  static void check(double a_, double k_) {
    if (!(a_ > 0.0) || !(k_ >= a_)) throw std::runtime_error("HuberLossWithCutoff: need a > 0 and k >= a");
  }
};

}  // namespace small_thing
