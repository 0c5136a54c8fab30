/*
 * Minimal base definitions of the dmlc namespace needed by the parser path.
 * API-compatible with the pieces of reference include/dmlc/base.h the path uses
 * (real_t / index_t live in data.h there; BeginPtr at base.h:237-257).
 */
#ifndef DMLC_BASE_H_
#define DMLC_BASE_H_

#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

#define DMLC_B200 1

namespace dmlc {

/* pointer to the first element, or NULL for an empty container (reference base.h:237-257) */
template <typename T>
inline T *BeginPtr(std::vector<T> &v) { return v.empty() ? NULL : &v[0]; }
template <typename T>
inline const T *BeginPtr(const std::vector<T> &v) { return v.empty() ? NULL : &v[0]; }
inline char *BeginPtr(std::string &s) { return s.empty() ? NULL : &s[0]; }
inline const char *BeginPtr(const std::string &s) { return s.empty() ? NULL : &s[0]; }

}  // namespace dmlc
#endif  // DMLC_BASE_H_
