// VisibilityOrderingB200.cpp -- defines BuildVisibilityOrderedGroup with the
// reference's signature (Common/VisibilityOrdering.hpp:14-23) on top of
// libamrvr_b200.so.  Compile it INSTEAD of Common/VisibilityOrdering.cpp.
//
// Same contract as the reference (Common/VisibilityOrdering.cpp:63-632):
// collective over `communicator`; returns a new MPI_Group the caller frees;
// never throws for a graph that cannot be ordered -- warns once on rank 0 and
// falls back to the identity order.  The box table travels over MPI exactly as
// there (:85-152); the depth-key sort and the pair test run in
// amrvr_order_records (device radix sort / pair kernel from 1024 boxes up).
// writeVisibilityGraph (:319-352): rank 0 writes visibility_graph_<n>.dot per
// round into the working directory from amrvr_order_graph_dot_host's text
// (byte-identical to the reference's, tests/test_cpu_host_logic.py).
#include <Common/VisibilityOrdering.hpp>

#include <cstdint>
#include <fstream>
#include <iostream>
#include <numeric>
#include <sstream>
#include <string>
#include <vector>

#include "B200Runtime.hpp"

MPI_Group BuildVisibilityOrderedGroup(
    const amrVolumeRenderer::volume::CameraParameters& camera,
    float aspect,
    MPI_Group baseGroup,
    int rank,
    int numProcs,
    bool useVisibilityGraph,
    bool writeVisibilityGraph,
    const std::vector<amrVolumeRenderer::volume::AmrBox>& localBoxes,
    MPI_Comm communicator) {
  std::vector<int> rankOrder(static_cast<std::size_t>(numProcs));
  std::iota(rankOrder.begin(), rankOrder.end(), 0);

  if (useVisibilityGraph && numProcs > 0) {
    // allgather of box counts, then of 6 corner values per box (rank-major)
    const int localCount = static_cast<int>(localBoxes.size());
    std::vector<int> counts(static_cast<std::size_t>(numProcs), 0);
    MPI_Allgather(&localCount, 1, MPI_INT, counts.data(), 1, MPI_INT, communicator);
    std::vector<int> valueCounts(counts.size()), displacements(counts.size());
    int total = 0;
    for (int r = 0; r < numProcs; ++r) {
      valueCounts[r] = counts[r] * 6;
      displacements[r] = total * 6;
      total += counts[r];
    }
    std::vector<double> mine(static_cast<std::size_t>(localCount) * 6);
    for (int b = 0; b < localCount; ++b)
      for (int d = 0; d < 3; ++d) {
        mine[6 * b + d] = localBoxes[b].minCorner[d];
        mine[6 * b + 3 + d] = localBoxes[b].maxCorner[d];
      }
    std::vector<double> all(static_cast<std::size_t>(total) * 6);
    MPI_Allgatherv(mine.data(), localCount * 6, MPI_DOUBLE, all.data(), valueCounts.data(),
                   displacements.data(), MPI_DOUBLE, communicator);

    std::vector<amrvr_box_record> records(static_cast<std::size_t>(total));
    int g = 0;
    for (int r = 0; r < numProcs; ++r)
      for (int l = 0; l < counts[r]; ++l, ++g) {
        amrvr_box_record& rec = records[g];
        rec = amrvr_box_record{};
        for (int d = 0; d < 3; ++d) {
          rec.min_corner[d] = all[6 * g + d];
          rec.max_corner[d] = all[6 * g + 3 + d];
          rec.dims[d] = 1;
        }
        rec.owner = r;
        rec.local_index = l;
      }
    int graphFailed = 0;
    amrvr_ctx* ctx = amrvr_b200::localContext(rank);
    const amrvr_camera abiCamera = amrvr_b200::toAbi(camera);
    amrvr_b200::check(ctx,
                      amrvr_order_records(ctx, numProcs, total, records.data(), &abiCamera, aspect,
                                          1, rankOrder.data(), &graphFailed),
                      "amrvr_order_records");
    if (writeVisibilityGraph && rank == 0) {
      static int graphFileCounter = 0;
      int rounds = 1;
      for (int round = 0; round < rounds; ++round) {
        std::int64_t bytes = 0;
        amrvr_b200::check(ctx,
                          amrvr_order_graph_dot_host(numProcs, total, records.data(), &abiCamera,
                                                     aspect, round, nullptr, 0, &bytes, &rounds),
                          "amrvr_order_graph_dot_host");
        if (round >= rounds) break;
        std::string text(static_cast<std::size_t>(bytes), '\0');
        amrvr_b200::check(ctx,
                          amrvr_order_graph_dot_host(numProcs, total, records.data(), &abiCamera,
                                                     aspect, round, text.data(), bytes, nullptr,
                                                     nullptr),
                          "amrvr_order_graph_dot_host");
        std::ostringstream filename;
        filename << "visibility_graph_" << graphFileCounter++ << ".dot";
        std::ofstream dotFile(filename.str(), std::ios::binary);
        if (!dotFile) {
          std::cerr << "Failed to write visibility graph to '" << filename.str() << "'"
                    << std::endl;
          continue;
        }
        dotFile << text;
        std::cout << "Wrote visibility graph to '" << filename.str() << "'" << std::endl;
      }
    }
    if (graphFailed != 0) {
      static bool warnedGraphFailure = false;
      if (!warnedGraphFailure && rank == 0) {
        std::cerr << "Visibility graph ordering failed; "
                     "falling back to default MPI rank order."
                  << std::endl;
      }
      warnedGraphFailure = true;
    }
  }

  MPI_Group orderedGroup = MPI_GROUP_NULL;
  MPI_Group_incl(baseGroup, numProcs, rankOrder.data(), &orderedGroup);
  return orderedGroup;
}
