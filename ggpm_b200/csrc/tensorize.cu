// Host-side index builder (no device code): batch-level padded index tensors of one graph family — what
// MolGraph.tensorize_graph builds with Python dict / list loops per edge (ggpm/mol_graph.py:238-281) — and the
// motif -> atom cluster table of MolGraph.tensorize (ggpm/mol_graph.py:211-222), from flat int32 arrays.
// SURVEY.md 8(f) rank 3.  Layout facts restated from the reference:
//   * ids are 1-based over the whole batch, id 0 = padding; row 0 of every table is the padding row;
//     fnode[0] duplicates fnode[1] (:276); one global edge counter across the batch (:265);
//   * edges are enumerated the way nx.DiGraph.edges does after convert_node_labels_to_integers: by source node in
//     node order, then in insertion order of that source's out-edges (:260);
//   * agraph[v] = ids of the edges INTO v, in enumeration order (:267);
//   * bgraph[e = (u -> v)] = ids of the edges (w -> u), w != v, in the order G.predecessors(u) yields them, which is
//     the enumeration order of the edges into u (:270-274);
//   * rows are zero-padded to (longest row + 1) columns (create_pad_tensor, ggpm/nnutils.py:105-110).
// Two calls: ggpm_tensorize_graph() builds into thread-local storage and reports the sizes, ggpm_tensorize_fetch()
// copies into caller-allocated arrays (so the caller owns all memory and widths are known before allocation).
#include <stdint.h>
#include <string.h>

#include <vector>

#include "common.cuh"

namespace ggpm {

struct TensorizeResult {
    int n_nodes1 = 0, n_mess1 = 0, label_cols = 0, a_width = 0, b_width = 0, n_mols = 0;
    std::vector<int32_t> fnode, fmess, scope;
    std::vector<int32_t> a_ptr, a_col, b_ptr, b_col;      // CSR of agraph / bgraph (rows incl. padding row 0)
};
static thread_local TensorizeResult g_tz;

}  // namespace ggpm

using namespace ggpm;

// node_off [n_mols+1], node_label [total_nodes, label_cols], edge_off [n_mols+1], edges [total_edges, 4] =
// (u, v, label, position) with u, v local to the molecule, in insertion order.
// sizes_out[6] = {rows of fnode (N+1), rows of fmess (E+1), agraph width, bgraph width, nnz(agraph), nnz(bgraph)}
extern "C" int ggpm_tensorize_graph(int n_mols, const int32_t* node_off, const int32_t* node_label, int label_cols,
                                    const int32_t* edge_off, const int32_t* edges, int32_t* sizes_out) {
    GGPM_REQUIRE(n_mols >= 1 && node_off && node_label && edge_off && edges && sizes_out, "tensorize_graph: bad arguments");
    GGPM_REQUIRE(label_cols >= 1 && label_cols <= 4, "tensorize_graph: label_cols must be 1..4");
    GGPM_REQUIRE(node_off[0] == 0 && edge_off[0] == 0, "tensorize_graph: offsets must start at 0");
    TensorizeResult& r = g_tz;
    const int N = node_off[n_mols], E = edge_off[n_mols];
    GGPM_REQUIRE(N >= 1, "tensorize_graph: no nodes");
    r.n_mols = n_mols; r.label_cols = label_cols; r.n_nodes1 = N + 1; r.n_mess1 = E + 1;
    r.fnode.assign((size_t)(N + 1) * label_cols, 0);
    r.fmess.assign((size_t)(E + 1) * 4, 0);
    r.scope.assign((size_t)n_mols * 2, 0);
    memcpy(r.fnode.data() + label_cols, node_label, sizeof(int32_t) * (size_t)N * label_cols);
    memcpy(r.fnode.data(), node_label, sizeof(int32_t) * label_cols);          // fnode[0] = fnode[1]

    std::vector<int32_t> src_cnt, order, into_ptr, into_src, into_eid, fill;
    r.a_ptr.assign(N + 2, 0);
    r.a_col.clear(); r.a_col.reserve(E);
    r.b_ptr.assign(E + 2, 0);
    r.b_col.clear(); r.b_col.reserve((size_t)E * 2);
    int eid = 0;                                       // global edge counter; ids start at 1
    for (int b = 0; b < n_mols; ++b) {
        const int n0 = node_off[b], n = node_off[b + 1] - n0;
        const int e0 = edge_off[b], m = edge_off[b + 1] - e0;
        GGPM_REQUIRE(n >= 0 && m >= 0, "tensorize_graph: offsets must be non-decreasing");
        const int offset = n0 + 1;
        r.scope[2 * b] = offset; r.scope[2 * b + 1] = n;
        const int32_t* ed = edges + (size_t)e0 * 4;
        // enumeration order: stable counting sort of the insertion-ordered edges by source
        src_cnt.assign(n + 1, 0);
        for (int j = 0; j < m; ++j) {
            const int u = ed[4 * j], v = ed[4 * j + 1];
            GGPM_REQUIRE(u >= 0 && u < n && v >= 0 && v < n, "tensorize_graph: molecule %d edge %d has node ids outside [0, %d)", b, j, n);
            src_cnt[u + 1]++;
        }
        for (int u = 0; u < n; ++u) src_cnt[u + 1] += src_cnt[u];
        order.resize(m);
        fill.assign(src_cnt.begin(), src_cnt.end() - 1);
        for (int j = 0; j < m; ++j) order[fill[ed[4 * j]]++] = j;
        // edges into every node, in enumeration order
        into_ptr.assign(n + 1, 0);
        for (int j = 0; j < m; ++j) into_ptr[ed[4 * j + 1] + 1]++;
        for (int v = 0; v < n; ++v) into_ptr[v + 1] += into_ptr[v];
        into_src.resize(m); into_eid.resize(m);
        fill.assign(into_ptr.begin(), into_ptr.end() - 1);
        const int first = eid + 1;
        for (int k = 0; k < m; ++k) {
            const int32_t* e = ed + 4 * (size_t)order[k];
            const int id = first + k;
            int32_t* fm = r.fmess.data() + (size_t)id * 4;
            fm[0] = e[0] + offset; fm[1] = e[1] + offset; fm[2] = e[2]; fm[3] = e[3];
            const int slot = fill[e[1]]++;
            into_src[slot] = e[0];
            into_eid[slot] = id;
        }
        // agraph rows of this molecule's nodes
        for (int v = 0; v < n; ++v) {
            for (int s = into_ptr[v]; s < into_ptr[v + 1]; ++s) r.a_col.push_back(into_eid[s]);
            r.a_ptr[offset + v + 1] = (int32_t)r.a_col.size();
        }
        // bgraph rows of this molecule's edges
        for (int k = 0; k < m; ++k) {
            const int32_t* e = ed + 4 * (size_t)order[k];
            const int u = e[0], v = e[1];
            for (int s = into_ptr[u]; s < into_ptr[u + 1]; ++s)
                if (into_src[s] != v) r.b_col.push_back(into_eid[s]);
            r.b_ptr[first + k + 1] = (int32_t)r.b_col.size();
        }
        eid += m;
    }
    r.a_ptr[1] = 0;                                    // padding rows are empty
    r.b_ptr[1] = 0;
    int aw = 0, bw = 0;
    for (int v = 0; v <= N; ++v) { const int l = r.a_ptr[v + 1] - r.a_ptr[v]; if (l > aw) aw = l; }
    for (int e = 0; e <= E; ++e) { const int l = r.b_ptr[e + 1] - r.b_ptr[e]; if (l > bw) bw = l; }
    r.a_width = aw + 1; r.b_width = bw + 1;
    sizes_out[0] = N + 1; sizes_out[1] = E + 1; sizes_out[2] = r.a_width; sizes_out[3] = r.b_width;
    sizes_out[4] = (int32_t)r.a_col.size(); sizes_out[5] = (int32_t)r.b_col.size();
    return 0;
}

// Copies the result of the last ggpm_tensorize_graph() of this thread.  Any output pointer may be null.
// fnode [N+1, label_cols], fmess [E+1, 4], agraph [N+1, a_width], bgraph [E+1, b_width] (zero padded), scope [n_mols, 2],
// and the CSR of both neighbour tables: a_ptr [N+2], a_col [nnz_a], b_ptr [E+2], b_col [nnz_b].
extern "C" int ggpm_tensorize_fetch(int32_t* fnode, int32_t* fmess, int32_t* agraph, int32_t* bgraph, int32_t* scope,
                                    int32_t* a_ptr, int32_t* a_col, int32_t* b_ptr, int32_t* b_col) {
    const TensorizeResult& r = g_tz;
    GGPM_REQUIRE(r.n_nodes1 > 0, "tensorize_fetch: no result on this thread");
    if (fnode) memcpy(fnode, r.fnode.data(), sizeof(int32_t) * r.fnode.size());
    if (fmess) memcpy(fmess, r.fmess.data(), sizeof(int32_t) * r.fmess.size());
    if (scope) memcpy(scope, r.scope.data(), sizeof(int32_t) * r.scope.size());
    if (agraph) {
        memset(agraph, 0, sizeof(int32_t) * (size_t)r.n_nodes1 * r.a_width);
        for (int v = 0; v < r.n_nodes1; ++v)
            memcpy(agraph + (size_t)v * r.a_width, r.a_col.data() + r.a_ptr[v], sizeof(int32_t) * (r.a_ptr[v + 1] - r.a_ptr[v]));
    }
    if (bgraph) {
        memset(bgraph, 0, sizeof(int32_t) * (size_t)r.n_mess1 * r.b_width);
        for (int e = 0; e < r.n_mess1; ++e)
            memcpy(bgraph + (size_t)e * r.b_width, r.b_col.data() + r.b_ptr[e], sizeof(int32_t) * (r.b_ptr[e + 1] - r.b_ptr[e]));
    }
    if (a_ptr) memcpy(a_ptr, r.a_ptr.data(), sizeof(int32_t) * (r.n_nodes1 + 1));
    if (a_col) memcpy(a_col, r.a_col.data(), sizeof(int32_t) * r.a_col.size());
    if (b_ptr) memcpy(b_ptr, r.b_ptr.data(), sizeof(int32_t) * (r.n_mess1 + 1));
    if (b_col) memcpy(b_col, r.b_col.data(), sizeof(int32_t) * r.b_col.size());
    return 0;
}

// cgraph [Nt+1, width]: atoms of every motif, shifted by the molecule's first atom id, zero padded — NO extra zero
// column (ggpm/mol_graph.py:211-222).  cluster_off [n_motifs+1] / cluster_atoms: atoms of motif i (local atom ids);
// motif_mol [n_motifs]: molecule of motif i; tree_scope / graph_scope [n_mols, 2].  Call once with cgraph == null to
// get the width in *width_out, then with the array.
extern "C" int ggpm_tensorize_cgraph(int n_mols, int n_motifs, const int32_t* cluster_off, const int32_t* cluster_atoms,
                                     const int32_t* motif_mol, const int32_t* tree_scope, const int32_t* graph_scope,
                                     int32_t* width_out, int32_t* cgraph) {
    GGPM_REQUIRE(n_mols >= 1 && n_motifs >= 1 && cluster_off && cluster_atoms && motif_mol && tree_scope && graph_scope && width_out,
                 "tensorize_cgraph: bad arguments");
    int w = 0;
    for (int i = 0; i < n_motifs; ++i) {
        const int l = cluster_off[i + 1] - cluster_off[i];
        GGPM_REQUIRE(l >= 0, "tensorize_cgraph: offsets must be non-decreasing");
        if (l > w) w = l;
    }
    *width_out = w;
    if (cgraph == nullptr) return 0;
    memset(cgraph, 0, sizeof(int32_t) * (size_t)(n_motifs + 1) * w);
    int b = 0, local = 0;                               // motifs arrive molecule by molecule
    for (int i = 0; i < n_motifs; ++i) {
        const int mol = motif_mol[i];
        GGPM_REQUIRE(mol >= 0 && mol < n_mols, "tensorize_cgraph: motif %d names molecule %d", i, mol);
        if (mol != b) { b = mol; local = 0; }
        const int row = tree_scope[2 * mol] + local++;
        GGPM_REQUIRE(row >= 1 && row <= n_motifs, "tensorize_cgraph: motif %d falls outside the tree scope", i);
        const int a0 = graph_scope[2 * mol];
        for (int k = cluster_off[i]; k < cluster_off[i + 1]; ++k)
            cgraph[(size_t)row * w + (k - cluster_off[i])] = cluster_atoms[k] + a0;
    }
    return 0;
}
