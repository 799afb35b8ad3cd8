// tree.h -- the tree side of the database build's host orchestration: what ipk::build is handed besides the posteriors.
//
// Restates, for the hot path's inputs only (i2l's phylo_tree / newick sources are absent):
//   newick reader / writer            i2l::io::load_newick, to_newick            (call sites ipk/src/extended_tree.cpp:167, db_builder.cpp:174)
//   ghost-node injection + mapping    tree_extender::extend_subtree              ipk/src/extended_tree.cpp:76-155 (branch lengths :7-73)
//   re-rooting of the AR tree         ipk::reroot_tree                           ipk/src/extended_tree.cpp:186-205
//   extended label -> AR label        ar::map_nodes                              ipk/src/ar.cpp:790-834
//   ghost selection and grouping      is_ghost, get_ghost_ids, group_ghost_ids   ipk/src/db_builder.cpp:495-553
//   tree index of the .ipk header     db_builder::run                            ipk/src/db_builder.cpp:191-197
#pragma once
#include <cstdint>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

namespace ipkhost
{
    struct pnode
    {
        std::string label;
        double branch_length = 0.0;
        int parent = -1;
        std::vector<int> children;
        // filled by ptree::index()
        uint32_t postorder_id = 0, preorder_id = 0;
        uint32_t num_nodes = 1, num_leaves = 1;      // nodes / leaves of the subtree rooted here (itself included)
        double subtree_branch_length = 0.0;          // sum of the branch lengths below this node (its own branch excluded) -- UNVERIFIED(i2l)
        bool is_leaf() const { return children.empty(); }
    };

    /// Rooted tree as the reference uses it: nodes addressed by post-order id, iteration = post-order.
    struct ptree
    {
        std::vector<pnode> nodes;
        int root = -1;
        std::vector<int> postorder;                  // node index by post-order id

        int add(const std::string& label, double len, int parent);
        void index();                                // post-/pre-order ids, subtree sizes and lengths
        size_t node_count() const { return nodes.size(); }
        bool is_rooted() const { return root >= 0 && nodes[root].children.size() < 3; }     // a trifurcating root = unrooted newick
    };

    ptree parse_newick(const std::string& text);     // throws std::runtime_error on malformed input
    std::string to_newick(const ptree& t);           // labels and branch lengths, children in stored order

    using ghost_mapping = std::unordered_map<std::string, uint32_t>;     // ghost label -> ORIGINAL post-order id
    /// extended_tree.cpp:86-150: every non-root node gets X0 on its branch, X1 under X0, dummy leaves X2, X3 under X1
    std::pair<ptree, ghost_mapping> extend_tree(const ptree& original);
    /// extended_tree.cpp:186-205: (a, b, c); -> ((b, c), a)added_root;
    void reroot_tree(ptree& t);
    /// ar.cpp:790-834: lock-step post-order walk; labelled extended nodes only
    std::unordered_map<std::string, std::string> map_nodes(const ptree& extended, const ptree& ar_tree);

    enum class ghost_strategy { inner_only, outer_only, both };
    struct ghost_plan
    {
        std::vector<std::string> ext_labels;         // ghost nodes in processing order: group after group, members in discovery order
        std::vector<uint32_t> node_branch;           // their ORIGINAL post-order id (the branch the k-mers are attributed to)
        uint32_t n_groups = 0;
    };
    /// db_builder.cpp:495-553: ghosts of the extended tree in post-order, filtered by strategy, grouped by original node in
    /// order of first appearance, the root's skipped
    ghost_plan plan_ghosts(const ptree& original, const ptree& extended, const ghost_mapping& mapping, ghost_strategy s);

    struct node_index { uint64_t num_nodes; double subtree_branch_length; };
    std::vector<node_index> tree_index(const ptree& original);            // db_builder.cpp:191-197, post-order
}
