// Throughput driver in the shape of the reference's test/libsvm_parser_test.cc:14-34
// (Create a parser for <file> partid npart, drain Next(), print MB/sec), written against the
// public API only so that it builds unchanged against this package's headers
// (dmlc_core_b200/include) or against the reference's own include/ directory.
//
//   g++ -std=c++14 -Idmlc_core_b200/include tests/cpp/libsvm_parser_driver.cc \
//       -Ldmlc_core_b200/lib -ldmlc_b200_host -ldmlc_b200 -o driver
//   ./driver <file> <partid> <npart> [type=libsvm] [passes=1]
#include <dmlc/data.h>
#include <dmlc/timer.h>

#include <cstdio>
#include <cstdlib>

int main(int argc, char *argv[]) {
  if (argc < 4) {
    std::printf("Usage: <file> partid npart [type]\n");
    return 0;
  }
  const char *type = argc > 4 ? argv[4] : "libsvm";
  dmlc::Parser<unsigned> *parser =
      dmlc::Parser<unsigned>::Create(argv[1], std::atoi(argv[2]), std::atoi(argv[3]), type);
  const int passes = argc > 5 ? std::atoi(argv[5]) : 1;   /* like csv_parser_test.cc: BeforeFirst */
  size_t before = 0;
  for (int pass = 0; pass < passes; ++pass) {
    if (pass != 0) parser->BeforeFirst();
    const double start = dmlc::GetTime();
    size_t rows = 0, nnz = 0;
    while (parser->Next()) {
      const dmlc::RowBlock<unsigned> &b = parser->Value();
      rows += b.size;
      nnz += b.offset[b.size] - b.offset[0];
    }
    const double dt = dmlc::GetTime() - start;
    const size_t bytes = parser->BytesRead() - before;
    before = parser->BytesRead();
    std::printf("%lu examples, %lu entries, %lu MB read, %g MB/sec\n", (unsigned long)rows,
                (unsigned long)nnz, (unsigned long)(bytes >> 20UL), (bytes >> 20UL) / dt);
  }
  delete parser;
  return 0;
}
