// Shim for the reference arm: reader/writer lock over pthread_rwlock.
#pragma once
#include <pthread.h>
namespace phi {
class RWLock {
 public:
  RWLock() { pthread_rwlock_init(&lock_, nullptr); }
  ~RWLock() { pthread_rwlock_destroy(&lock_); }
  void RDLock() { pthread_rwlock_rdlock(&lock_); }
  void WRLock() { pthread_rwlock_wrlock(&lock_); }
  void UNLock() { pthread_rwlock_unlock(&lock_); }
  RWLock(const RWLock&) = delete;
  RWLock& operator=(const RWLock&) = delete;
 private:
  pthread_rwlock_t lock_;
};
}  // namespace phi
