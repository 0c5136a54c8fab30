/*
 * Public data API of the parser path, source-compatible with reference include/dmlc/data.h:
 * real_t / index_t (:23,:29), DataIter (:52-63), Row (:69-158), RowBlock (:169-231),
 * RowBlockIter (:246-267), Parser (:283-311), ParserFactoryReg (:317-320) and
 * DMLC_REGISTER_DATA_PARSER (:347-350).  Parser<IndexType>::Create returns the B200 parser
 * (dmlc_core_b200/src/b200_text_parser.cc) for the registered types libsvm, libfm and csv.
 */
#ifndef DMLC_DATA_H_
#define DMLC_DATA_H_

#include <map>
#include <string>

#include "./base.h"
#include "./io.h"
#include "./logging.h"
#include "./registry.h"

namespace dmlc {

typedef float real_t;      /* label / weight / value type */
typedef unsigned index_t;  /* default feature index type */

/* forward iterator over values of DType */
template <typename DType>
class DataIter {
 public:
  virtual ~DataIter(void) {}
  virtual void BeforeFirst(void) = 0;
  virtual bool Next(void) = 0;
  virtual const DType &Value(void) const = 0;
};

/* one sparse row: a view into a RowBlock */
template <typename IndexType>
class Row {
 public:
  const real_t *label;
  const real_t *weight;     /* NULL: 1.0 */
  const uint64_t *qid;      /* NULL: 0 */
  size_t length;
  const IndexType *field;   /* NULL unless LibFM */
  const IndexType *index;
  const real_t *value;      /* NULL: every value is 1.0 */

  inline IndexType get_field(size_t i) const { return field[i]; }
  inline IndexType get_index(size_t i) const { return index[i]; }
  inline real_t get_value(size_t i) const { return value == NULL ? 1.0f : value[i]; }
  inline real_t get_label() const { return *label; }
  inline real_t get_weight() const { return weight == NULL ? 1.0f : *weight; }
  inline uint64_t get_qid() const { return qid == NULL ? 0 : *qid; }
  /* sparse dot product with a dense vector of `size` entries */
  template <typename V>
  inline V SDot(const V *dense, size_t size) const {
    V acc = static_cast<V>(0);
    for (size_t i = 0; i < length; ++i) {
      CHECK(index[i] < size) << "feature index exceed bound";
      acc += value == NULL ? dense[index[i]] : dense[index[i]] * value[i];
    }
    return acc;
  }
};

/* a batch of rows in CSR layout; arrays that do not exist are NULL */
template <typename IndexType>
struct RowBlock {
  size_t size;            /* number of rows */
  const size_t *offset;   /* [size + 1] */
  const real_t *label;    /* [size] */
  const real_t *weight;
  const uint64_t *qid;
  const IndexType *field;
  const IndexType *index;
  const real_t *value;

  inline Row<IndexType> operator[](size_t rowid) const {
    CHECK(rowid < size);
    Row<IndexType> r;
    const size_t at = offset[rowid];
    r.label = label + rowid;
    r.weight = weight == NULL ? NULL : weight + rowid;
    r.qid = qid == NULL ? NULL : qid + rowid;
    r.length = offset[rowid + 1] - at;
    r.field = field == NULL ? NULL : field + at;
    r.index = index + at;
    r.value = value == NULL ? NULL : value + at;
    return r;
  }
  inline size_t MemCostBytes(void) const {
    const size_t nnz = offset[size] - offset[0];
    size_t cost = size * (sizeof(size_t) + sizeof(real_t));
    if (weight != NULL) cost += size * sizeof(real_t);
    if (qid != NULL) cost += size * sizeof(size_t);
    if (field != NULL) cost += nnz * sizeof(IndexType);
    if (index != NULL) cost += nnz * sizeof(IndexType);
    if (value != NULL) cost += nnz * sizeof(real_t);
    return cost;
  }
  /* rows [begin, end); offset keeps indexing the parent's arrays */
  inline RowBlock Slice(size_t begin, size_t end) const {
    CHECK(begin <= end && end <= size);
    RowBlock part = *this;
    part.size = end - begin;
    part.offset = offset + begin;
    part.label = label + begin;
    part.weight = weight == NULL ? NULL : weight + begin;
    part.qid = qid == NULL ? NULL : qid + begin;
    return part;
  }
};

/* iterator over the RowBlocks of a whole (partition of a) data set */
template <typename IndexType>
class RowBlockIter : public DataIter<RowBlock<IndexType> > {
 public:
  /* uri may carry ?format=..&label_column=.. arguments; type "auto" reads format= */
  static RowBlockIter<IndexType> *Create(const char *uri, unsigned part_index,
                                         unsigned num_parts, const char *type);
  /* 1 + the largest feature index seen */
  virtual size_t NumCol() const = 0;
};

/* streaming parser: one RowBlock per Next() */
template <typename IndexType>
class Parser : public DataIter<RowBlock<IndexType> > {
 public:
  static Parser<IndexType> *Create(const char *uri_, unsigned part_index, unsigned num_parts,
                                   const char *type);
  /* bytes of text consumed so far */
  virtual size_t BytesRead(void) const = 0;
  /* plugin hook: builds a parser for a path + URI arguments */
  typedef Parser<IndexType> *(*Factory)(const std::string &path,
                                        const std::map<std::string, std::string> &args,
                                        unsigned part_index, unsigned num_parts);
};

template <typename IndexType>
struct ParserFactoryReg
    : public FunctionRegEntryBase<ParserFactoryReg<IndexType>, typename Parser<IndexType>::Factory> {};

}  // namespace dmlc

/* DMLC_REGISTER_DATA_PARSER(uint32_t, mytype, MyFactory<uint32_t>); */
#define DMLC_REGISTER_DATA_PARSER(IndexType, TypeName, FactoryFunction)                       \
  DMLC_REGISTRY_REGISTER(::dmlc::ParserFactoryReg<IndexType>, ParserFactoryReg##_##IndexType, \
                         TypeName)                                                            \
      .set_body(FactoryFunction)

#endif  // DMLC_DATA_H_
