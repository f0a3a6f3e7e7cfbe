// Driver-facing data model of the k-NN engine.
//
// API parity with the reference driver contract (reference: common.h:4-27):
// the four aggregate types below keep the reference's names and field order so
// that user code written against `Params` / `DataPoint` / `Query` / `Update`
// and the `reportResult` callback compiles unchanged.  Everything the B200
// engine uses internally is structure-of-arrays (see include/dmlp/views.h);
// these array-of-structs types exist only at the API boundary.
#ifndef DMLP_COMMON_H_
#define DMLP_COMMON_H_

#include <utility>
#include <vector>

/// Problem header: first line of the input stream (`N Q A`).
struct Params {
  int num_data;     ///< N: number of labelled points
  int num_queries;  ///< Q: number of queries
  int num_attrs;    ///< A: feature dimension
};

/// One labelled point; `id` is its 0-based line index in the input.
struct DataPoint {
  int id;
  int label;
  std::vector<double> attrs;
};

/// One query; `id` is its 0-based index, `k` the neighbour count it asks for.
struct Query {
  int id;
  int k;
  std::vector<double> attrs;
};

/// Point update record.  Kept for API parity; the reference never uses it
/// (reference: common.h:22-25, common.cpp:46-55 is dead code) and neither do we.
struct Update {
  int id;
  std::vector<double> new_attrs;
};

/// Result sink.  `result` holds (distance, datapoint id) pairs already ordered
/// by ascending distance, ties broken by larger id first.  Release builds print
/// `Query <id> checksum: <u64>`; -DDEBUG builds print the label and neighbours
/// (reference: common.cpp:57-79).
void reportResult(Query &q, std::vector<std::pair<double, int>> &result, int label);

#endif  // DMLP_COMMON_H_
