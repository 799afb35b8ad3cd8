// plan.cpp -- C entry points over tree.h: the build plan (ipkhost.h).
#include "ipkhost.h"
#include "tree.h"

#include <stdexcept>
#include <string>

namespace ipkhost { void set_last_error(const std::string& s); }

struct ipkhost_plan
{
    ipkhost::ghost_plan ghosts;
    std::vector<std::string> ar_labels;
    std::vector<uint64_t> index_num_nodes;
    std::vector<double> index_lengths;
    std::string newick, extended_newick;
    uint32_t node_count = 0;
};

extern "C"
{

int ipkhost_plan_create(const char* original_newick, const char* ar_tree_newick, int strategy, int use_unrooted, ipkhost_plan** out)
{
    if (!original_newick || !out || strategy < 0 || strategy > 2) return IPKHOST_EINVAL;
    *out = nullptr;
    try
    {
        using namespace ipkhost;
        const ptree original = parse_newick(original_newick);
        if (!original.is_rooted() && !use_unrooted)
        {
            set_last_error("This reference tree is not rooted. Please provide a rooted tree or provide --use-unrooted. "
                           "WARNING! This may impact placement accuracy.");
            return IPKHOST_EINVAL;
        }
        auto [extended, mapping] = extend_tree(original);
        auto plan = new ipkhost_plan();
        const ghost_strategy s = strategy == 1 ? ghost_strategy::inner_only : strategy == 2 ? ghost_strategy::outer_only : ghost_strategy::both;
        plan->ghosts = plan_ghosts(original, extended, mapping, s);
        if (ar_tree_newick)
        {
            ptree ar_tree = parse_newick(ar_tree_newick);
            if (original.is_rooted() && !ar_tree.is_rooted()) reroot_tree(ar_tree);        // main.cpp:172-178
            std::unordered_map<std::string, std::string> m;
            try { m = map_nodes(extended, ar_tree); }
            catch (const std::exception& e) { delete plan; set_last_error(e.what()); return IPKHOST_EFORMAT; }
            for (const auto& l : plan->ghosts.ext_labels)
            {
                const auto it = m.find(l);
                if (it == m.end()) { delete plan; set_last_error("no AR node for " + l); return IPKHOST_ENOENT; }
                plan->ar_labels.push_back(it->second);
            }
        }
        else plan->ar_labels = plan->ghosts.ext_labels;
        for (const auto& ni : tree_index(original)) { plan->index_num_nodes.push_back(ni.num_nodes); plan->index_lengths.push_back(ni.subtree_branch_length); }
        plan->newick = to_newick(original);
        plan->extended_newick = to_newick(extended);
        plan->node_count = uint32_t(original.node_count());
        *out = plan;
        return IPKHOST_OK;
    }
    catch (const std::exception& e)
    {
        ipkhost::set_last_error(e.what());
        return IPKHOST_EFORMAT;
    }
}

void ipkhost_plan_free(ipkhost_plan* p) { delete p; }
uint32_t ipkhost_plan_num_ghosts(const ipkhost_plan* p) { return p ? uint32_t(p->ghosts.ext_labels.size()) : 0; }
uint32_t ipkhost_plan_num_groups(const ipkhost_plan* p) { return p ? p->ghosts.n_groups : 0; }
const char* ipkhost_plan_ext_label(const ipkhost_plan* p, uint32_t i) { return p && i < p->ghosts.ext_labels.size() ? p->ghosts.ext_labels[i].c_str() : nullptr; }
const char* ipkhost_plan_ar_label(const ipkhost_plan* p, uint32_t i) { return p && i < p->ar_labels.size() ? p->ar_labels[i].c_str() : nullptr; }
const uint32_t* ipkhost_plan_node_branch(const ipkhost_plan* p) { return p ? p->ghosts.node_branch.data() : nullptr; }
uint32_t ipkhost_plan_node_count(const ipkhost_plan* p) { return p ? p->node_count : 0; }
const uint64_t* ipkhost_plan_index_num_nodes(const ipkhost_plan* p) { return p ? p->index_num_nodes.data() : nullptr; }
const double* ipkhost_plan_index_lengths(const ipkhost_plan* p) { return p ? p->index_lengths.data() : nullptr; }
const char* ipkhost_plan_newick(const ipkhost_plan* p) { return p ? p->newick.c_str() : nullptr; }
const char* ipkhost_plan_extended_newick(const ipkhost_plan* p) { return p ? p->extended_newick.c_str() : nullptr; }

}   // extern "C"
