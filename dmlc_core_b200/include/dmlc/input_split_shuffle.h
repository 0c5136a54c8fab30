/*
 * InputSplitShuffle for the stand-alone host layer: a partition read as `num_shuffle_parts`
 * sub-partitions in an order that is reshuffled on every BeforeFirst().
 * Same interface and same observable order as the reference's class
 * (reference include/dmlc/input_split_shuffle.h:19-165): part k of n with s sub-parts reads the
 * parts  order[j] + k * s  of  n * s  of the underlying split, j = 0 .. s-1, where `order` is
 * 0 .. s-1 shuffled by std::shuffle with a std::mt19937 seeded
 * 666 + part_index + num_parts + num_shuffle_parts + shuffle_seed (:113-118), and shuffled
 * again, with the same engine, by each BeforeFirst() (:24-30).  ResetPartition re-aims at
 * another part without reshuffling (:76-81).  With one sub-part it is the plain split.
 */
#ifndef DMLC_INPUT_SPLIT_SHUFFLE_H_
#define DMLC_INPUT_SPLIT_SHUFFLE_H_

#include <algorithm>
#include <memory>
#include <random>
#include <vector>

#include "./io.h"

namespace dmlc {

class InputSplitShuffle : public InputSplit {
 public:
  InputSplitShuffle(const char *uri, unsigned part_index, unsigned num_parts, const char *type,
                    unsigned num_shuffle_parts, int shuffle_seed)
      : part_(part_index), parts_(num_parts), subs_(num_shuffle_parts), at_(0) {
    order_.resize(subs_);
    for (unsigned i = 0; i < subs_; ++i) order_[i] = static_cast<int>(i);
    engine_.seed(kSeedBase + part_ + parts_ + subs_ + shuffle_seed);
    std::shuffle(order_.begin(), order_.end(), engine_);
    inner_.reset(InputSplit::Create(uri, SubPart(part_, 0), parts_ * subs_, type));
  }
  static InputSplit *Create(const char *uri, unsigned part_index, unsigned num_parts,
                            const char *type, unsigned num_shuffle_parts, int shuffle_seed) {
    CHECK(num_shuffle_parts > 0) << "number of shuffle parts should be greater than zero!";
    return new InputSplitShuffle(uri, part_index, num_parts, type, num_shuffle_parts, shuffle_seed);
  }

  virtual void HintChunkSize(size_t chunk_size) { inner_->HintChunkSize(chunk_size); }
  virtual size_t GetTotalSize(void) { return inner_->GetTotalSize(); }
  virtual void BeforeFirst(void) {
    if (subs_ <= 1) {
      inner_->BeforeFirst();
      return;
    }
    std::shuffle(order_.begin(), order_.end(), engine_);
    at_ = 0;
    inner_->ResetPartition(SubPart(part_, 0), parts_ * subs_);
  }
  virtual bool NextRecord(Blob *out_rec) {
    while (!inner_->NextRecord(out_rec)) {
      if (!NextSubPart()) return false;
    }
    return true;
  }
  virtual bool NextChunk(Blob *out_chunk) {
    while (!inner_->NextChunk(out_chunk)) {
      if (!NextSubPart()) return false;
    }
    return true;
  }
  virtual void ResetPartition(unsigned part_index, unsigned num_parts) {
    CHECK(num_parts == parts_) << "num_parts is not consistent!";
    at_ = 0;
    inner_->ResetPartition(SubPart(part_index, 0), num_parts * subs_);
  }

 private:
  static const int kSeedBase = 666;
  unsigned SubPart(unsigned part, unsigned j) const {
    return static_cast<unsigned>(order_[j]) + part * subs_;
  }
  /* move on to the next sub-partition of this part; false when all have been read */
  bool NextSubPart() {
    if (subs_ <= 1 || at_ + 1 >= subs_) return false;
    ++at_;
    inner_->ResetPartition(SubPart(part_, at_), parts_ * subs_);
    return true;
  }
  std::mt19937 engine_;
  std::unique_ptr<InputSplit> inner_;
  unsigned part_, parts_, subs_, at_;
  std::vector<int> order_;
};

}  // namespace dmlc
#endif  // DMLC_INPUT_SPLIT_SHUFFLE_H_
