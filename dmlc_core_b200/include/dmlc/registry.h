/*
 * Name -> factory registry, API-compatible with the subset of reference
 * include/dmlc/registry.h the parser plugin hook needs: Registry<E>::Get(), Find(), List(),
 * __REGISTER__ (rejects duplicate names, registry.h:78-80), FunctionRegEntryBase with
 * set_body / describe, DMLC_REGISTRY_ENABLE and DMLC_REGISTRY_REGISTER (registry.h:230-248).
 */
#ifndef DMLC_REGISTRY_H_
#define DMLC_REGISTRY_H_

#include <map>
#include <string>
#include <vector>

#include "./logging.h"

namespace dmlc {

template <typename EntryType>
class Registry {
 public:
  /* singleton, defined once per EntryType by DMLC_REGISTRY_ENABLE */
  static Registry *Get();

  const EntryType *Find(const std::string &name) const {
    typename std::map<std::string, EntryType *>::const_iterator it = by_name_.find(name);
    return it == by_name_.end() ? NULL : it->second;
  }
  const std::vector<const EntryType *> &List() const { return const_list_; }
  std::vector<std::string> ListAllNames() const {
    std::vector<std::string> names;
    for (typename std::map<std::string, EntryType *>::const_iterator it = by_name_.begin();
         it != by_name_.end(); ++it) names.push_back(it->first);
    return names;
  }
  /* registers a new entry; a second registration of the same name is fatal */
  EntryType &__REGISTER__(const std::string &name) {
    CHECK_EQ(by_name_.count(name), 0U) << name << " already registered";
    EntryType *e = new EntryType();
    e->name = name;
    by_name_[name] = e;
    const_list_.push_back(e);
    return *e;
  }
  EntryType &__REGISTER_OR_GET__(const std::string &name) {
    if (by_name_.count(name) != 0) return *by_name_[name];
    return __REGISTER__(name);
  }

 private:
  Registry() {}
  ~Registry() {
    for (size_t i = 0; i < const_list_.size(); ++i) delete const_list_[i];
  }
  std::map<std::string, EntryType *> by_name_;
  std::vector<const EntryType *> const_list_;
};

/* common base of function-like registry entries */
template <typename EntryType, typename FunctionType>
class FunctionRegEntryBase {
 public:
  std::string name;
  std::string description;
  FunctionType body;
  std::string return_type;

  EntryType &set_body(FunctionType f) {
    body = f;
    return static_cast<EntryType &>(*this);
  }
  EntryType &describe(const std::string &text) {
    description = text;
    return static_cast<EntryType &>(*this);
  }
  EntryType &set_return_type(const std::string &type) {
    return_type = type;
    return static_cast<EntryType &>(*this);
  }
};

}  // namespace dmlc

#define DMLC_REGISTRY_ENABLE(EntryType)                \
  template <>                                          \
  Registry<EntryType> *Registry<EntryType>::Get() {    \
    static Registry<EntryType> inst;                   \
    return &inst;                                      \
  }

#define DMLC_STR_CONCAT_(a, b) a##b
#define DMLC_STR_CONCAT(a, b) DMLC_STR_CONCAT_(a, b)
#define DMLC_ATTRIBUTE_UNUSED __attribute__((unused))

#define DMLC_REGISTRY_REGISTER(EntryType, EntryTypeName, Name)                   \
  static DMLC_ATTRIBUTE_UNUSED EntryType &__make_##EntryTypeName##_##Name##__ =  \
      ::dmlc::Registry<EntryType>::Get()->__REGISTER__(#Name)

#endif  // DMLC_REGISTRY_H_
