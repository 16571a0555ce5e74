// Compile-and-link check of the C++ operator API: user code written the way the reference's own examples and tests call
// it (cpp/src/examples/join_example.cpp, cpp/test/{join,set_op,groupby,partition,table_op}_test.cpp) must build against
// include/cylon/*.hpp and link against libcylon_b200.so. main() returns before touching a GPU unless asked to run.
#include <cstdio>
#include <cstring>
#include <memory>
#include <unordered_map>
#include <vector>

#include <cylon/ctx/cylon_context.hpp>
#include <cylon/groupby/groupby.hpp>
#include <cylon/join/join_config.hpp>
#include <cylon/net/comm_config.hpp>
#include <cylon/partition/partition.hpp>
#include <cylon/status.hpp>
#include <cylon/table.hpp>

static int run() {
  auto ctx = cylon::CylonContext::Init();
  std::shared_ptr<cylon::Table> left, right, out;
  std::vector<cyl_column_t> host_cols;  // Arrow buffers of host arrays go here
  std::vector<std::string> names;
  cylon::Status st = cylon::Table::FromHostColumns(ctx, host_cols, names, left);
  if (!st.is_ok()) return st.get_code();
  st = cylon::Table::FromHostColumns(ctx, host_cols, names, right);

  // joins: table.hpp:258,269 and join/join_config.hpp:33-189
  auto inner = cylon::join::config::JoinConfig::InnerJoin(0, 0, cylon::join::config::JoinAlgorithm::HASH);
  auto left_join = cylon::join::config::JoinConfig::LeftJoin(std::vector<int>{0, 1}, std::vector<int>{0, 1},
                                                             cylon::join::config::JoinAlgorithm::SORT, "l_", "r_");
  auto right_join = cylon::join::config::JoinConfig::RightJoin(0, 0);
  auto full = cylon::join::config::JoinConfig::FullOuterJoin(0, 0);
  st = cylon::Join(left, right, inner, out);
  st = cylon::DistributedJoin(left, right, left_join, out);
  (void)right_join;
  (void)full;

  // set operators, unique, shuffle, hash partition: table.hpp:279-348,433-436
  st = cylon::Union(left, right, out);
  st = cylon::Subtract(left, right, out);
  st = cylon::Intersect(left, right, out);
  st = cylon::DistributedUnion(left, right, out);
  st = cylon::DistributedSubtract(left, right, out);
  st = cylon::DistributedIntersect(left, right, out);
  st = cylon::Unique(left, {0}, out, true);
  st = cylon::DistributedUnique(left, {0}, out);
  st = cylon::Shuffle(left, {0}, out);
  std::unordered_map<int, std::shared_ptr<cylon::Table>> parts;
  st = cylon::HashPartition(left, {0}, 4, &parts);

  // group-by: groupby/groupby.hpp:23-33, groupby/hash_groupby.hpp:31-66
  st = cylon::DistributedHashGroupBy(left, 0, {1, 1}, {cylon::compute::SUM, cylon::compute::COUNT}, out);
  st = cylon::HashGroupBy(left, std::vector<int32_t>{0}, {{1, cylon::compute::MIN}, {1, cylon::compute::MAX}}, out);

  // partition + split: partition/partition.hpp:32-85
  uint32_t *pid = nullptr;
  uint64_t *hist = nullptr;
  st = cylon::MapToHashPartitions(left, 0, 8, &pid, &hist);
  std::vector<std::shared_ptr<cylon::Table>> pieces;
  st = cylon::Split(left, 8, pid, pieces);

  // table.hpp:353-404 and partition/partition.hpp:60-74 as cpp/test/partition_test.cpp:81-153 calls them
  st = cylon::Sort(left, 0, out);
  st = cylon::Sort(left, 0, out, false);
  st = cylon::Sort(left, std::vector<int32_t>{0, 1}, out, true);
  st = cylon::Sort(left, std::vector<int32_t>{0, 1}, out, std::vector<bool>{true, false});
  st = cylon::DistributedSort(left, 1, out);
  st = cylon::DistributedSort(left, 0, out, true, {(uint32_t)4, (uint64_t)left->Rows()});
  st = cylon::DistributedSort(left, {0, 1}, out, {true, false}, {(uint32_t)4, (uint64_t)left->Rows()});
  std::vector<uint32_t> sort_partitions, sort_count;
  st = cylon::MapToSortPartitions(left, 0, 4, sort_partitions, sort_count, true, left->Rows(), 4);
  st = cylon::Split(left, 4, sort_partitions, sort_count, pieces);

  std::printf("%lld rows, %d columns, rank %d of %d, retain %d\n", (long long)out->Rows(), out->Columns(), ctx->GetRank(),
              ctx->GetWorldSize(), (int)out->IsRetain());
  out->retainMemory(false);
  ctx->Barrier();
  ctx->Finalize();
  return st.is_ok() ? 0 : st.get_code();
}

int main(int argc, char **argv) {
  if (argc > 1 && std::strcmp(argv[1], "run") == 0) return run();
  return 0;  // building and linking is the test
}
