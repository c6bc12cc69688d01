// genogrove_b200/chain_shape.hpp — counts-only twin of one index's B+ tree.
//
// The device answers every query from the LEAF-CHAIN ORDER of an index, and almost every insert path fixes a
// key's place in that order by its value alone.  The one exception is the sorted single append: it runs
// lower_bound inside the RIGHTMOST LEAF only (grove_insert.ipp:501-531, node.hpp:247-254), so where a key lands
// among EQUAL keys depends on how many keys that leaf holds — and that depends on every split, borrow and merge
// the tree went through.  None of those decisions look at key values: splits cut at fixed midpoints
// (grove.hpp:284-286, grove_insert.ipp:316), removal rebalances by key COUNTS (grove_remove.ipp:153-273), and the
// leaf a regular insert or a removal touches is the leaf that holds the key's chain rank (routing on subtree maxima,
// grove_insert.ipp:236-244, picks the first child whose last key is not below the new key, i.e. the child that
// covers the key's lower_bound rank).  So the tree's SHAPE can be replayed from ranks alone.  chain_shape keeps
// exactly that — per node a child list or a key count, plus subtree totals for rank navigation — and applies the
// reference's structural rules to it; the facade asks it for the rightmost leaf's key count.
//
// Rules restated (reference file:line):
//   regular insert     descend to the child covering the rank, an empty child taking everything (grove_insert.ipp:
//                      236-244); on the way back split any child that reached `order` keys, leaves and internal nodes at
//                      mid = order / 2 (:259-261,306-318,344-391); a root with `order` keys gets a new root above (:190-194)
//   sorted append      +1 key in the rightmost leaf; while a node on the rightmost spine has `order` keys split it — a
//                      leaf at order - 1, leaving ONE key right, internal nodes at order / 2 (:457-473,316)
//   sorted bulk        empty index: leaves hold distribute_evenly(n, order - 1) keys, each upper layer
//                      distribute_evenly(children, order) (grove.hpp:95-106, grove_insert.ipp:548-647); otherwise one
//                      sorted append per key (:106-126)
//   removal            -1 key in the leaf holding the rank; a non-root leaf below ceil((order-1)/2) keys — internal node
//                      below floor((order-1)/2) — borrows one from its left sibling if that has more than the minimum,
//                      else from the right one, else merges with the left sibling (the right one for a first child);
//                      the parent is then collapsed (a root left with one child), rebalanced the same way or left alone
//                      (grove_remove.ipp:24-62,153-273; grove.hpp:272-279)
#ifndef GENOGROVE_B200_CHAIN_SHAPE_HPP
#define GENOGROVE_B200_CHAIN_SHAPE_HPP

#include <algorithm>
#include <cstddef>
#include <cstdint>
#include <deque>
#include <memory>
#include <span>
#include <stdexcept>
#include <vector>

namespace genogrove_b200::structure::detail {

class chain_shape {
  public:
    /// pre-order token of a leaf: bit 31 set, low bits = key count; of an internal node: its child count
    static constexpr std::uint32_t leaf_bit = 0x80000000u;

    explicit chain_shape(int order = 3) : order_(order), pool_(std::make_unique<std::deque<node>>()) {}

    [[nodiscard]] bool empty() const noexcept { return root_ == nullptr; }
    [[nodiscard]] std::size_t size() const noexcept { return root_ != nullptr ? root_->total : 0; }
    [[nodiscard]] int order() const noexcept { return order_; }

    void clear() {
        pool_->clear();
        free_.clear();
        root_ = nullptr;
    }

    /// key count of the last leaf of the chain (0 for an empty index)
    [[nodiscard]] std::size_t rightmost_leaf_keys() const noexcept {
        const node* n = root_;
        if (n == nullptr) return 0;
        while (!n->leaf) n = n->kids.back();
        return n->nkeys;
    }

    /// bottom-up build of n > 0 keys into an empty shape
    void build_bottom_up(std::size_t n) {
        clear();
        if (n == 0) return;
        std::vector<node*> layer;
        const even leaves = distribute(n, static_cast<std::size_t>(order_ - 1));
        layer.reserve(leaves.groups);
        for (std::size_t i = 0; i < leaves.groups; ++i) {
            node* l = make(true);
            l->nkeys = static_cast<std::uint32_t>(leaves.count_for(i));
            l->total = l->nkeys;
            layer.push_back(l);
        }
        while (layer.size() > 1) {
            std::vector<node*> parents;
            const even d = distribute(layer.size(), static_cast<std::size_t>(order_));
            parents.reserve(d.groups);
            std::size_t c = 0;
            for (std::size_t p = 0; p < d.groups; ++p) {
                node* par = make(false);
                for (std::size_t i = 0; i < d.count_for(p); ++i, ++c) adopt(par, layer[c]);
                parents.push_back(par);
            }
            layer.swap(parents);
        }
        root_ = layer[0];
    }

    /// rebuild from pre-order tokens (what a serialized grove's node blocks give, gg_loader.hpp)
    void build_from_tokens(std::span<const std::uint32_t> tokens) {
        clear();
        if (tokens.empty()) return;
        std::size_t at = 0;
        root_ = parse(tokens, at, 0);
        if (at != tokens.size()) throw std::invalid_argument("chain_shape: trailing tokens");
    }

    /// pre-order tokens of the current shape
    [[nodiscard]] std::vector<std::uint32_t> tokens() const {
        std::vector<std::uint32_t> out;
        if (root_ != nullptr) dump(root_, out);
        return out;
    }

    /// a regular insert whose key lands at chain rank `rank` (its lower_bound over the whole chain)
    void insert_at(std::size_t rank) {
        if (root_ == nullptr) root_ = make(true);
        node* n = root_;
        while (!n->leaf) {
            const std::size_t last = n->kids.size() - 1;
            std::size_t ci = 0, before = 0;
            while (ci < last) {
                const node* c = n->kids[ci];
                if (c->total == 0 || rank < before + c->total) break;
                before += c->total;
                ++ci;
            }
            rank -= before;
            n = n->kids[ci];
        }
        ++n->nkeys;
        for (node* p = n; p != nullptr; p = p->parent) ++p->total;
        for (node* child = n; child->parent != nullptr;) {
            node* par = child->parent;
            if (key_count(child) == order_) split_child(par, position(child), false);
            child = par;
        }
        if (key_count(root_) == order_) promote_new_root(false);
    }

    /// a sorted append (single, or one key of a sorted bulk onto a non-empty index)
    void append_sorted() {
        if (root_ == nullptr) {
            root_ = make(true);
            root_->nkeys = 1;
            root_->total = 1;
            return;
        }
        node* r = root_;
        while (!r->leaf) r = r->kids.back();
        ++r->nkeys;
        for (node* p = r; p != nullptr; p = p->parent) ++p->total;
        for (node* over = r; key_count(over) == order_;) {
            if (over->parent == nullptr) {
                promote_new_root(true);
                break;
            }
            node* par = over->parent;
            split_child(par, par->kids.size() - 1, true);
            over = par;
        }
    }

    /// removal of the key at chain rank `rank`
    void remove_at(std::size_t rank) {
        if (root_ == nullptr || rank >= root_->total) throw std::out_of_range("chain_shape::remove_at");
        node* n = root_;
        while (!n->leaf) {
            std::size_t ci = 0;
            while (rank >= n->kids[ci]->total) rank -= n->kids[ci++]->total;
            n = n->kids[ci];
        }
        --n->nkeys;
        for (node* p = n; p != nullptr; p = p->parent) --p->total;
        if (n->parent == nullptr) {
            if (n->nkeys == 0) clear();
            return;
        }
        if (key_count(n) < min_keys(n)) rebalance(n);
    }

  private:
    struct node {
        node* parent = nullptr;
        std::vector<node*> kids;  // internal nodes; keys = kids.size() - 1
        std::size_t total = 0;    // leaf keys beneath
        std::uint32_t nkeys = 0;  // leaves
        bool leaf = true;
    };
    struct even {
        std::size_t groups, base, extra;
        [[nodiscard]] std::size_t count_for(std::size_t i) const noexcept { return base + (i < extra ? 1 : 0); }
    };
    static even distribute(std::size_t total, std::size_t max_per_group) noexcept {
        const std::size_t groups = total / max_per_group + (total % max_per_group != 0 ? 1 : 0);
        return {groups, total / groups, total % groups};
    }

    node* make(bool leaf) {
        node* n;
        if (!free_.empty()) {
            n = free_.back();
            free_.pop_back();
            *n = node{};
        } else {
            n = &pool_->emplace_back();
        }
        n->leaf = leaf;
        return n;
    }
    void release(node* n) {
        n->kids.clear();
        free_.push_back(n);
    }
    static void adopt(node* par, node* child) {
        par->kids.push_back(child);
        child->parent = par;
        par->total += child->total;
    }
    [[nodiscard]] int key_count(const node* n) const noexcept {
        return n->leaf ? static_cast<int>(n->nkeys) : static_cast<int>(n->kids.size()) - 1;
    }
    [[nodiscard]] int min_keys(const node* n) const noexcept { return n->leaf ? order_ / 2 : (order_ - 1) / 2; }  // ceil / floor of (order-1)/2
    static std::size_t position(const node* n) {
        const auto& sib = n->parent->kids;
        return static_cast<std::size_t>(std::find(sib.begin(), sib.end(), n) - sib.begin());
    }
    static std::size_t sum(const node* n) {
        std::size_t t = 0;
        for (const node* c : n->kids) t += c->total;
        return t;
    }

    void split_child(node* par, std::size_t idx, bool sorted_append) {
        node* child = par->kids[idx];
        node* sib = make(child->leaf);
        sib->parent = par;
        if (child->leaf) {
            const std::uint32_t mid = static_cast<std::uint32_t>(sorted_append ? order_ - 1 : order_ / 2);
            sib->nkeys = child->nkeys - mid;
            child->nkeys = mid;
            sib->total = sib->nkeys;
            child->total = child->nkeys;
        } else {
            const std::size_t mid = static_cast<std::size_t>(order_ / 2);  // left keeps children [0, mid], key[mid] moves up
            sib->kids.assign(child->kids.begin() + static_cast<std::ptrdiff_t>(mid) + 1, child->kids.end());
            child->kids.resize(mid + 1);
            for (node* c : sib->kids) c->parent = sib;
            sib->total = sum(sib);
            child->total = sum(child);
        }
        par->kids.insert(par->kids.begin() + static_cast<std::ptrdiff_t>(idx) + 1, sib);
    }

    void promote_new_root(bool sorted_append) {
        node* top = make(false);
        adopt(top, root_);
        root_ = top;
        split_child(top, 0, sorted_append);
    }

    void rebalance(node* n) {
        node* par = n->parent;
        const std::size_t pos = position(n);
        if (pos > 0) {  // borrow from the left sibling
            node* left = par->kids[pos - 1];
            if (key_count(left) > min_keys(n)) {
                if (n->leaf) {
                    --left->nkeys, --left->total;
                    ++n->nkeys, ++n->total;
                } else {
                    node* moved = left->kids.back();
                    left->kids.pop_back();
                    left->total -= moved->total;
                    n->kids.insert(n->kids.begin(), moved);
                    moved->parent = n;
                    n->total += moved->total;
                }
                return;
            }
        }
        if (pos + 1 < par->kids.size()) {  // borrow from the right sibling
            node* right = par->kids[pos + 1];
            if (key_count(right) > min_keys(n)) {
                if (n->leaf) {
                    --right->nkeys, --right->total;
                    ++n->nkeys, ++n->total;
                } else {
                    node* moved = right->kids.front();
                    right->kids.erase(right->kids.begin());
                    right->total -= moved->total;
                    n->kids.push_back(moved);
                    moved->parent = n;
                    n->total += moved->total;
                }
                return;
            }
        }
        // merge: into the left sibling, or the right sibling into a first child
        node* left = pos > 0 ? par->kids[pos - 1] : n;
        node* right = pos > 0 ? n : par->kids[pos + 1];
        const std::size_t right_pos = pos > 0 ? pos : pos + 1;
        if (left->leaf) {
            left->nkeys += right->nkeys;
        } else {
            for (node* c : right->kids) c->parent = left;
            left->kids.insert(left->kids.end(), right->kids.begin(), right->kids.end());
        }
        left->total += right->total;
        par->kids.erase(par->kids.begin() + static_cast<std::ptrdiff_t>(right_pos));
        release(right);
        if (par == root_ && par->kids.size() == 1) {
            root_ = par->kids[0];
            root_->parent = nullptr;
            release(par);
        } else if (par != root_ && key_count(par) < min_keys(par)) {
            rebalance(par);
        }
    }

    node* parse(std::span<const std::uint32_t> tokens, std::size_t& at, int depth) {
        if (at >= tokens.size() || depth > 64) throw std::invalid_argument("chain_shape: malformed tokens");
        const std::uint32_t t = tokens[at++];
        if ((t & leaf_bit) != 0) {
            node* l = make(true);
            l->nkeys = t & ~leaf_bit;
            l->total = l->nkeys;
            return l;
        }
        if (t == 0) throw std::invalid_argument("chain_shape: internal node without children");
        node* n = make(false);
        n->kids.reserve(t);
        for (std::uint32_t i = 0; i < t; ++i) adopt(n, parse(tokens, at, depth + 1));
        return n;
    }
    static void dump(const node* n, std::vector<std::uint32_t>& out) {
        if (n->leaf) {
            out.push_back(leaf_bit | n->nkeys);
            return;
        }
        out.push_back(static_cast<std::uint32_t>(n->kids.size()));
        for (const node* c : n->kids) dump(c, out);
    }

    int order_;
    std::unique_ptr<std::deque<node>> pool_;  // stable addresses; behind a pointer so that moving the shape keeps them
    std::vector<node*> free_;
    node* root_ = nullptr;
};

}  // namespace genogrove_b200::structure::detail

#endif  // GENOGROVE_B200_CHAIN_SHAPE_HPP
