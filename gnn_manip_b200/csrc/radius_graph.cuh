#pragma once
#include "common.cuh"

namespace gm {

struct RadiusGraphWs {
  int max_cells;
  unsigned* bbox;
  char* grid;
  int* cell_of;
  int* cell_cnt;
  int* cell_start;
  float4* sorted;
  int* indeg;
  int* fill;
  int* keys;
  int* scan_ws;
};

size_t radius_graph_workspace_bytes(int n_total, int n_graphs, int K);
int radius_graph_max_cells(int n_total);
void radius_graph_carve(Arena& a, int n_total, int n_graphs, int K, RadiusGraphWs& ws);

// pos: first coordinate of node 0; consecutive nodes are pos_stride floats apart (x,y,z contiguous).
int radius_graph(const float* pos, int pos_stride, int n_per_graph, int n_graphs, double r, int K, int* nbr,
                 int* cnt, int* edge_ptr, int* recv_ptr, int* esend, int* erecv, int* eref, int* n_edges,
                 const RadiusGraphWs& ws, cudaStream_t stream);

// Receiver CSR over edge slots.  K > 0: slot = sender*K + s (neighbour table, -1 = empty);
// K == 0: slot = edge id of an explicit edge list (send_of_slot gives the sender).
int build_recv_csr(const int* recv_of_slot, long long n_slots, const int* n_valid_dev, int n_nodes, int K,
                   const int* edge_ptr, const int* send_of_slot, int* indeg, int* fill, int* keys, int* scan_ws,
                   bool indeg_ready, int* recv_ptr, int* esend, int* erecv, int* eref, cudaStream_t stream,
                   bool fill_ready = false, bool recv_ptr_ready = false);

}  // namespace gm
