// ORACLE / test infrastructure: single-threaded stand-in for the two boost names the reference's data classes use.
#pragma once
namespace boost {
class mutex {
 public:
  class scoped_lock { public: explicit scoped_lock(mutex&) {} };
};
}  // namespace boost
