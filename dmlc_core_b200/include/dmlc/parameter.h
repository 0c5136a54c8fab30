/*
 * Only what the parser path needs from reference include/dmlc/parameter.h: the exception type
 * raised for bad URI arguments (parameter.h:30). The CSV factory validates its two arguments
 * (format, label_column) by hand instead of through the generic Parameter<> machinery.
 */
#ifndef DMLC_PARAMETER_H_
#define DMLC_PARAMETER_H_
#include <string>
#include "./logging.h"
namespace dmlc {
struct ParamError : public dmlc::Error {
  explicit ParamError(const std::string &msg) : dmlc::Error(msg) {}
};
}  // namespace dmlc
#endif  // DMLC_PARAMETER_H_
