// Stand-in for the reference's metrics interfaces (cartographer/metrics/{counter,gauge,histogram,
// family_factory}.h) so that the constraint builders keep reporting what the reference reports
// (constraint_builder_2d.cc:46-53,318-343).  In a full Cartographer tree these headers exist
// already and this file is not needed; here they also come with an in-memory factory that the
// tests and the bench read back.
#ifndef CARTOGRAPHER_METRICS_FAMILY_FACTORY_H_
#define CARTOGRAPHER_METRICS_FAMILY_FACTORY_H_

#include <algorithm>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

namespace cartographer {
namespace metrics {

class Counter {
 public:
  static Counter* Null();
  virtual ~Counter() = default;
  virtual void Increment() = 0;
  virtual void Increment(double by_value) = 0;
};

class Gauge {
 public:
  static Gauge* Null();
  virtual ~Gauge() = default;
  virtual void Increment() = 0;
  virtual void Increment(double by_value) = 0;
  virtual void Decrement() = 0;
  virtual void Decrement(double by_value) = 0;
  virtual void Set(double value) = 0;
};

class Histogram {
 public:
  using BucketBoundaries = std::vector<double>;
  static Histogram* Null();
  static BucketBoundaries FixedWidth(double width, int num_finite_buckets) {
    BucketBoundaries b;
    double upper = width;
    for (int i = 0; i < num_finite_buckets; ++i, upper += width) b.push_back(upper);
    return b;
  }
  virtual ~Histogram() = default;
  virtual void Observe(double value) = 0;
};

namespace detail {
struct NullCounter : Counter {
  void Increment() override {}
  void Increment(double) override {}
};
struct NullGauge : Gauge {
  void Increment() override {}
  void Increment(double) override {}
  void Decrement() override {}
  void Decrement(double) override {}
  void Set(double) override {}
};
struct NullHistogram : Histogram {
  void Observe(double) override {}
};
}  // namespace detail

inline Counter* Counter::Null() {
  static detail::NullCounter c;
  return &c;
}
inline Gauge* Gauge::Null() {
  static detail::NullGauge g;
  return &g;
}
inline Histogram* Histogram::Null() {
  static detail::NullHistogram h;
  return &h;
}

template <typename MetricType>
class Family {
 public:
  virtual ~Family() = default;
  virtual MetricType* Add(const std::map<std::string, std::string>& labels) = 0;
};

class FamilyFactory {
 public:
  virtual ~FamilyFactory() = default;
  virtual Family<Counter>* NewCounterFamily(const std::string& name, const std::string& description) = 0;
  virtual Family<Gauge>* NewGaugeFamily(const std::string& name, const std::string& description) = 0;
  virtual Family<Histogram>* NewHistogramFamily(const std::string& name, const std::string& description,
                                                const Histogram::BucketBoundaries& boundaries) = 0;
};

// --- addition of this repository: a factory that keeps its metrics in memory -------------------
// Keys are "family{label=value,...}"; counters and gauges read back as one double, histograms as
// the per-bucket counts (one more bucket than boundaries: the overflow bucket).
class InMemoryFamilyFactory : public FamilyFactory {
 public:
  Family<Counter>* NewCounterFamily(const std::string& name, const std::string&) override {
    return Keep(new FamilyOf<Counter, Scalar<Counter>>(this, name, {}));
  }
  Family<Gauge>* NewGaugeFamily(const std::string& name, const std::string&) override {
    return Keep(new FamilyOf<Gauge, Scalar<Gauge>>(this, name, {}));
  }
  Family<Histogram>* NewHistogramFamily(const std::string& name, const std::string&,
                                        const Histogram::BucketBoundaries& boundaries) override {
    return Keep(new FamilyOf<Histogram, Buckets>(this, name, boundaries));
  }
  double Value(const std::string& key) {
    std::lock_guard<std::mutex> l(mu_);
    auto it = scalars_.find(key);
    return it == scalars_.end() ? 0. : *it->second;
  }
  std::vector<double> BucketCounts(const std::string& key) {
    std::lock_guard<std::mutex> l(mu_);
    auto it = buckets_.find(key);
    return it == buckets_.end() ? std::vector<double>() : *it->second;
  }

 private:
  template <typename Base>
  struct Scalar : Base {
    Scalar(InMemoryFamilyFactory* f, const std::string& key, const Histogram::BucketBoundaries&)
        : f(f), v(f->ScalarSlot(key)) {}
    void Add(double by) {
      std::lock_guard<std::mutex> l(f->mu_);
      *v += by;
    }
    void Increment() { Add(1.); }
    void Increment(double by) { Add(by); }
    void Decrement() { Add(-1.); }
    void Decrement(double by) { Add(-by); }
    void Set(double value) {
      std::lock_guard<std::mutex> l(f->mu_);
      *v = value;
    }
    InMemoryFamilyFactory* f;
    double* v;
  };
  struct Buckets : Histogram {
    Buckets(InMemoryFamilyFactory* f, const std::string& key, const BucketBoundaries& b)
        : f(f), c(f->BucketSlot(key)), b(b) {
      c->assign(b.size() + 1, 0.);
    }
    void Observe(double value) override {
      std::lock_guard<std::mutex> l(f->mu_);
      (*c)[std::lower_bound(b.begin(), b.end(), value) - b.begin()] += 1.;
    }
    InMemoryFamilyFactory* f;
    std::vector<double>* c;
    BucketBoundaries b;
  };
  struct AnyFamily {
    virtual ~AnyFamily() = default;
  };
  template <typename Metric, typename Impl>
  struct FamilyOf : Family<Metric>, AnyFamily {
    FamilyOf(InMemoryFamilyFactory* f, const std::string& name, const Histogram::BucketBoundaries& b)
        : f(f), name(name), boundaries(b) {}
    Metric* Add(const std::map<std::string, std::string>& labels) override {
      std::string key = name + "{";
      for (const auto& kv : labels) key += (key.back() == '{' ? "" : ",") + kv.first + "=" + kv.second;
      key += "}";
      metrics.emplace_back(new Impl(f, key, boundaries));
      return metrics.back().get();
    }
    InMemoryFamilyFactory* f;
    std::string name;
    Histogram::BucketBoundaries boundaries;
    std::vector<std::unique_ptr<Impl>> metrics;
  };
  double* ScalarSlot(const std::string& key) {
    std::lock_guard<std::mutex> l(mu_);
    auto& slot = scalars_[key];
    if (!slot) slot.reset(new double(0.));
    return slot.get();
  }
  std::vector<double>* BucketSlot(const std::string& key) {
    std::lock_guard<std::mutex> l(mu_);
    auto& slot = buckets_[key];
    if (!slot) slot.reset(new std::vector<double>());
    return slot.get();
  }
  template <typename F>
  F* Keep(F* family) {
    families_.emplace_back(family);
    return family;
  }
  std::mutex mu_;
  std::map<std::string, std::unique_ptr<double>> scalars_;
  std::map<std::string, std::unique_ptr<std::vector<double>>> buckets_;
  std::vector<std::unique_ptr<AnyFamily>> families_;
};

}  // namespace metrics
}  // namespace cartographer

#endif  // CARTOGRAPHER_METRICS_FAMILY_FACTORY_H_
