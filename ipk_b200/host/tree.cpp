// tree.cpp -- see tree.h.  Host-only, O(nodes); nothing here is on the timed path.
#include "tree.h"

#include <cctype>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <cstdio>
#include <stdexcept>

namespace ipkhost
{
    int ptree::add(const std::string& label, double len, int parent)
    {
        pnode n;
        n.label = label; n.branch_length = len; n.parent = parent;
        nodes.push_back(std::move(n));
        const int id = int(nodes.size()) - 1;
        if (parent >= 0) nodes[parent].children.push_back(id);
        return id;
    }

    void ptree::index()
    {
        postorder.clear();
        postorder.reserve(nodes.size());
        if (root < 0) return;
        // iterative DFS (10,000-taxon caterpillar trees would overflow a recursive one)
        uint32_t pre = 0;
        std::vector<std::pair<int, size_t>> stack;
        stack.push_back({ root, 0 });
        nodes[root].preorder_id = pre++;
        while (!stack.empty())
        {
            auto& [v, next] = stack.back();
            if (next < nodes[v].children.size())
            {
                const int c = nodes[v].children[next++];
                nodes[c].preorder_id = pre++;
                stack.push_back({ c, 0 });
                continue;
            }
            pnode& n = nodes[v];
            n.postorder_id = uint32_t(postorder.size());
            postorder.push_back(v);
            n.num_nodes = 1; n.num_leaves = n.children.empty() ? 1 : 0; n.subtree_branch_length = 0.0;
            for (int c : n.children)
            {
                n.num_nodes += nodes[c].num_nodes;
                n.num_leaves += nodes[c].num_leaves;
                n.subtree_branch_length += nodes[c].subtree_branch_length + nodes[c].branch_length;
            }
            stack.pop_back();
        }
    }

    // ---- newick ------------------------------------------------------------------------------------------------------
    namespace
    {
        struct parser
        {
            const std::string& s;
            size_t i = 0;
            ptree t;
            explicit parser(const std::string& text) : s(text) {}
            [[noreturn]] void fail(const char* what) const
            {
                throw std::runtime_error(std::string("newick: ") + what + " at offset " + std::to_string(i));
            }
            void skip()
            {
                while (i < s.size())
                {
                    if (isspace(static_cast<unsigned char>(s[i]))) { ++i; continue; }
                    if (s[i] == '[') { while (i < s.size() && s[i] != ']') ++i; if (i < s.size()) ++i; continue; }      // comment
                    break;
                }
            }
            std::string label()
            {
                skip();
                std::string out;
                if (i < s.size() && (s[i] == '\'' || s[i] == '"'))
                {
                    const char q = s[i++];
                    while (i < s.size())
                    {
                        if (s[i] == q) { if (i + 1 < s.size() && s[i + 1] == q) { out += q; i += 2; continue; } ++i; break; }
                        out += s[i++];
                    }
                    return out;
                }
                while (i < s.size() && !strchr("(),:;[", s[i]) && !isspace(static_cast<unsigned char>(s[i]))) out += s[i++];
                return out;
            }
            double length()
            {
                skip();
                if (i >= s.size() || s[i] != ':') return 0.0;
                ++i;
                skip();
                const char* b = s.c_str() + i;
                char* e = nullptr;
                const double v = strtod(b, &e);
                if (e == b) fail("branch length expected");
                i += size_t(e - b);
                return v;
            }
            void run()
            {
                // iterative: a stack of open inner nodes
                std::vector<int> open;
                int cur = -1;                              // the node whose label / length come next
                skip();
                if (i >= s.size()) fail("empty input");
                bool done = false;
                while (!done)
                {
                    skip();
                    if (i >= s.size()) fail("unexpected end (missing ';')");
                    const char c = s[i];
                    if (c == '(')
                    {
                        ++i;
                        const int v = t.add("", 0.0, open.empty() ? -1 : open.back());
                        if (open.empty()) { if (t.root >= 0) fail("second root"); t.root = v; }
                        open.push_back(v);
                        cur = -1;
                        // a leaf may follow directly
                        skip();
                        if (i < s.size() && s[i] != '(')
                        {
                            const int leaf = t.add("", 0.0, open.back());
                            t.nodes[leaf].label = label();
                            t.nodes[leaf].branch_length = length();
                            cur = leaf;
                        }
                    }
                    else if (c == ',')
                    {
                        ++i;
                        if (open.empty()) fail("',' outside parentheses");
                        skip();
                        if (i < s.size() && s[i] != '(')
                        {
                            const int leaf = t.add("", 0.0, open.back());
                            t.nodes[leaf].label = label();
                            t.nodes[leaf].branch_length = length();
                            cur = leaf;
                        }
                    }
                    else if (c == ')')
                    {
                        ++i;
                        if (open.empty()) fail("unbalanced ')'");
                        cur = open.back();
                        open.pop_back();
                        t.nodes[cur].label = label();
                        t.nodes[cur].branch_length = length();
                    }
                    else if (c == ';')
                    {
                        ++i;
                        done = true;
                    }
                    else
                    {
                        // a tree that is a single leaf: "A;"
                        if (t.root >= 0 || !open.empty()) fail("unexpected character");
                        t.root = t.add("", 0.0, -1);
                        t.nodes[t.root].label = label();
                        t.nodes[t.root].branch_length = length();
                    }
                }
                if (!open.empty()) fail("unbalanced '('");
                if (t.root < 0) fail("no tree");
                (void)cur;
                t.index();
            }
        };

        bool needs_quotes(const std::string& l)
        {
            for (char c : l) if (strchr("(),:;[]' \t\n", c)) return true;
            return false;
        }
    }

    ptree parse_newick(const std::string& text)
    {
        parser p(text);
        p.run();
        return std::move(p.t);
    }

    std::string to_newick(const ptree& t)
    {
        // (snprintf, not iostreams: this library may be linked with a static libstdc++ and loaded next to another copy,
        // whose locale facets do not mix)
        std::string os;
        struct frame { int v; size_t next; };
        std::vector<frame> stack;
        if (t.root < 0) return ";";
        stack.push_back({ t.root, 0 });
        while (!stack.empty())
        {
            frame& f = stack.back();
            const pnode& n = t.nodes[f.v];
            if (f.next == 0 && !n.children.empty()) os += '(';
            if (f.next < n.children.size())
            {
                if (f.next) os += ',';
                const int c = n.children[f.next++];
                stack.push_back({ c, 0 });
                continue;
            }
            if (!n.children.empty()) os += ')';
            if (needs_quotes(n.label)) { os += '\''; os += n.label; os += '\''; } else os += n.label;
            if (f.v != t.root)
            {
                char buf[40];
                snprintf(buf, sizeof buf, ":%.10g", n.branch_length);
                os += buf;
            }
            stack.pop_back();
        }
        os += ';';
        return os;
    }

    // ---- ghost nodes -------------------------------------------------------------------------------------------------
    namespace
    {
        /// extended_tree.cpp:7-32 on the ORIGINAL tree
        double total_branch_length(const ptree& t, int root)
        {
            if (t.nodes[root].is_leaf()) return 0.0;
            double length = 0.0;
            // visit_subtree(root): every node of the subtree, the root included
            std::vector<int> stack{ root };
            while (!stack.empty())
            {
                const int v = stack.back();
                stack.pop_back();
                const pnode& n = t.nodes[v];
                length += n.is_leaf() ? n.branch_length : double(n.num_leaves) * n.branch_length;
                for (int c : n.children) stack.push_back(c);
            }
            length -= double(t.nodes[root].num_leaves) * t.nodes[root].branch_length;
            return length;
        }

        /// extended_tree.cpp:35-73: lengths of (x0 -> parent) and (x1 -> x0)
        std::pair<double, double> ghost_branch_lengths(const ptree& t, int node)
        {
            const pnode& n = t.nodes[node];
            const double old_len = n.branch_length;
            const double x0 = old_len / 2.0;
            const double residual = old_len - x0;
            double x1;
            if (n.is_leaf()) x1 = residual;
            else x1 = (total_branch_length(t, node) + residual * double(n.num_leaves)) / double(n.num_leaves);
            return { x0, x1 };
        }
    }

    std::pair<ptree, ghost_mapping> extend_tree(const ptree& original)
    {
        ptree ext = original;                       // node indices coincide with the original's until ghosts are appended
        ghost_mapping mapping;
        size_t counter = original.node_count() + 1;                                  // extended_tree.cpp:81
        // extend_subtree recurses into the children first: original post-order
        for (int v : original.postorder)
        {
            if (v == original.root) continue;
            const int parent = ext.nodes[v].parent;
            const auto [x0_len, x1_len] = ghost_branch_lengths(original, v);
            const std::string x0_name = std::to_string(counter++) + "_X0";
            // parent->remove_child(node); parent->add_child(x0): x0 goes to the END of the parent's list
            auto& pc = ext.nodes[parent].children;
            for (size_t i = 0; i < pc.size(); ++i) if (pc[i] == v) { pc.erase(pc.begin() + long(i)); break; }
            const int x0 = ext.add(x0_name, x0_len, parent);
            const std::string x1_name = std::to_string(counter++) + "_X1";
            const int x1 = ext.add(x1_name, x1_len, x0);
            ext.nodes[x0].children.push_back(v);                                     // x0: [x1, node]
            ext.nodes[v].parent = x0;
            ext.nodes[v].branch_length = ext.nodes[v].branch_length - x0_len;
            ext.add(std::to_string(counter++) + "_X2", 0.01, x1);
            ext.add(std::to_string(counter++) + "_X3", 0.01, x1);
            mapping[x0_name] = original.nodes[v].postorder_id;
            mapping[x1_name] = original.nodes[v].postorder_id;
        }
        ext.index();
        return { std::move(ext), std::move(mapping) };
    }

    void reroot_tree(ptree& t)
    {
        if (t.root < 0) return;
        if (t.nodes[t.root].children.size() > 2)
        {
            const int old_root = t.root;
            const int a = t.nodes[old_root].children[0];
            const int added = t.add("added_root", 0.0, -1);
            t.nodes[added].children.push_back(old_root);
            t.nodes[old_root].parent = added;
            t.nodes[added].children.push_back(a);
            t.nodes[a].parent = added;
            auto& rc = t.nodes[old_root].children;
            rc.erase(rc.begin());
            t.root = added;
            t.index();
        }
    }

    std::unordered_map<std::string, std::string> map_nodes(const ptree& extended, const ptree& ar_tree)
    {
        if (extended.node_count() != ar_tree.node_count())
            throw std::runtime_error("Error during database construction: extended tree and AR differ in the number of nodes: " +
                                     std::to_string(extended.node_count()) + " vs. " + std::to_string(ar_tree.node_count()));
        std::unordered_map<std::string, std::string> ext_to_ar;
        for (size_t i = 0; i < extended.postorder.size(); ++i)
        {
            const pnode& e = extended.nodes[extended.postorder[i]];
            if (e.label.empty()) continue;
            ext_to_ar[e.label] = ar_tree.nodes[ar_tree.postorder[i]].label;
        }
        return ext_to_ar;
    }

    namespace
    {
        bool ends_with(const std::string& s, const char* suffix)
        {
            const size_t n = strlen(suffix);
            return s.size() >= n && s.compare(s.size() - n, n, suffix) == 0;
        }
    }

    ghost_plan plan_ghosts(const ptree& original, const ptree& extended, const ghost_mapping& mapping, ghost_strategy s)
    {
        std::vector<std::vector<std::string>> groups;
        std::vector<uint32_t> group_branch;
        std::unordered_map<uint32_t, size_t> index;
        for (int v : extended.postorder)
        {
            const std::string& label = extended.nodes[v].label;
            const bool x0 = ends_with(label, "_X0"), x1 = ends_with(label, "_X1");
            const bool ghost = s == ghost_strategy::inner_only ? x0 : s == ghost_strategy::outer_only ? x1 : (x0 || x1);     // db_builder.cpp:495-507
            if (!ghost) continue;
            const auto it = mapping.find(label);
            if (it == mapping.end()) throw std::runtime_error("ghost node " + label + " is not in the ghost mapping");
            const uint32_t original_id = it->second;
            if (original_id < original.postorder.size() && original.postorder[original_id] == original.root) continue;      // :535-539
            const auto g = index.find(original_id);
            if (g != index.end()) groups[g->second].push_back(label);
            else
            {
                index[original_id] = groups.size();
                groups.push_back({ label });
                group_branch.push_back(original_id);
            }
        }
        ghost_plan p;
        p.n_groups = uint32_t(groups.size());
        for (size_t g = 0; g < groups.size(); ++g)
            for (auto& l : groups[g]) { p.ext_labels.push_back(l); p.node_branch.push_back(group_branch[g]); }
        return p;
    }

    std::vector<node_index> tree_index(const ptree& original)
    {
        std::vector<node_index> out;
        out.reserve(original.node_count());
        for (int v : original.postorder) out.push_back({ original.nodes[v].num_nodes, original.nodes[v].subtree_branch_length });
        return out;
    }
}
