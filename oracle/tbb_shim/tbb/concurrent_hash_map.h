// TEST INFRASTRUCTURE ONLY — stand-in for tbb::concurrent_hash_map (see concurrent_vector.h): one
// mutex around a std::unordered_map, with the accessor protocol HashBlock.hpp:43-101 uses
// (find / insert with accessor, size, begin / end).  Node addresses are stable, as TBB's are.
#pragma once
#include <memory>
#include <mutex>
#include <unordered_map>
namespace tbb {
template <typename Key, typename T, typename HashCompare>
class concurrent_hash_map {
  struct Hash {
    std::size_t operator()(const Key &k) const { return HashCompare::hash(k); }
  };
  struct Eq {
    bool operator()(const Key &a, const Key &b) const { return HashCompare::equal(a, b); }
  };
  using Map = std::unordered_map<Key, T, Hash, Eq>;

 public:
  using value_type = typename Map::value_type;
  class const_accessor {
   public:
    const value_type *operator->() const { return p_; }
    const value_type &operator*() const { return *p_; }
    bool empty() const { return !p_; }
   protected:
    friend class concurrent_hash_map;
    value_type *p_ = nullptr;
  };
  class accessor : public const_accessor {
   public:
    value_type *operator->() const { return this->p_; }
    value_type &operator*() const { return *this->p_; }
  };
  bool find(const_accessor &a, const Key &k) const {
    std::lock_guard<std::mutex> g(mu_);
    auto it = m_.find(k);
    a.p_ = it == m_.end() ? nullptr : const_cast<value_type *>(&*it);
    return a.p_ != nullptr;
  }
  bool find(accessor &a, const Key &k) {
    std::lock_guard<std::mutex> g(mu_);
    auto it = m_.find(k);
    a.p_ = it == m_.end() ? nullptr : &*it;
    return a.p_ != nullptr;
  }
  bool insert(accessor &a, const Key &k) {
    std::lock_guard<std::mutex> g(mu_);
    auto r = m_.try_emplace(k);
    a.p_ = &*r.first;
    return r.second;
  }
  std::size_t size() const { return m_.size(); }
  auto begin() { return m_.begin(); }
  auto end() { return m_.end(); }

 private:
  mutable std::mutex mu_;
  Map m_;
};
}  // namespace tbb
