// Shim for the reference arm: Place / CUDAPlace / CPUPlace carry a kind and a device id.
#pragma once
namespace paddle {
namespace platform {
class Place {
 public:
  enum Kind { kCPU = 0, kGPU = 1 };
  Place() : kind_(kCPU), dev_(0) {}
  Place(Kind k, int d) : kind_(k), dev_(d) {}
  int GetDeviceId() const { return dev_; }
  int device() const { return dev_; }
  Kind kind() const { return kind_; }
  bool operator==(const Place& o) const { return kind_ == o.kind_ && dev_ == o.dev_; }
 private:
  Kind kind_;
  int dev_;
};
class CUDAPlace : public Place {
 public:
  CUDAPlace() : Place(kGPU, 0) {}
  explicit CUDAPlace(int d) : Place(kGPU, d) {}
};
class CPUPlace : public Place {
 public:
  CPUPlace() : Place(kCPU, 0) {}
};
inline bool is_gpu_place(const Place& p) { return p.kind() == Place::kGPU; }
inline bool is_cpu_place(const Place& p) { return p.kind() == Place::kCPU; }
}  // namespace platform
}  // namespace paddle
