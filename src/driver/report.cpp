// Result sink shared by every backend (reference: common.cpp:57-79).
#include <cstdio>
#include <iostream>

#include "common.h"
#include "dmlp/views.h"

void reportResult(Query &q, std::vector<std::pair<double, int>> &result, int label) {
#ifndef DEBUG
  // Fold the label, then every neighbour id in reporting order.
  std::vector<int> ids;
  ids.reserve(result.size());
  for (const auto &entry : result) ids.push_back(entry.second);
  unsigned long long h = dmlp::result_checksum(label, ids.data(), static_cast<long long>(ids.size()));
  std::cout << "Query " << q.id << " checksum: " << h << "\n";
#else
  std::cout << "Label for Query " << q.id << " : " << label << "\n";
  std::cout << "Top-" << q.k << " neighbors:\n";
  for (const auto &entry : result) std::cout << entry.second << " : " << entry.first << "\n";
#endif
}
