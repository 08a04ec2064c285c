// .panman output for the MSA construction path: the tree built by pmfs_host::Tree serialised the way the reference's CLI
// does it (paths relative to /root/reference):
//   writePanMAN(globalVm, TG)            src/panmanUtils.cpp:269-297   xz/LZMA level 9 around the Cap'n Proto stream
//   TreeGroup::writeToFile               src/panman.cpp:7027-7161      TreeGroup{trees[Tree{...}], complexMutations[]}
//   Tree::getNodesPreorder               src/panman.cpp:2868-2943      nodes in pre-order; per node one Mutation per block
//                                                                       with nucMutation[{nucPosition, nucGapPosition,
//                                                                       nucGapExist, mutInfo = (nucs >> (24-len*4)) << 8 | mutInfo}]
//   Tree::getNewickString / splitNewick  src/panman.cpp:1923-2030, :2633-2644   Newick text in 1 MiB chunks (:44)
//   schema                               panman.capnp
// Cap'n Proto and liblzma are not on this image, so the message is encoded by hand: the unpacked wire format is a few
// rules (struct = data words + pointer words; struct pointer = offset | sizes; list pointer = offset | element size |
// count; struct lists carry a tag word), written here for exactly the structs of panman.capnp that this path fills.
// One segment, standard stream framing -- what capnp::writeMessage produces for a message that fits one segment; any
// Cap'n Proto reader accepts it.  The container layer is /usr/bin/xz -9 (the .xz format boost's lzma_compressor writes).
#include <cstdio>
#include <cstring>
#include <fstream>

#include "msa_tree.hpp"

namespace pmfs_host {

// ---- Newick text of the tree (src/panman.cpp:1923-2030), iterative: same characters, no recursion --------------------
std::string getNewickString(const Tree& T) {
    const std::vector<Node>& N = T.nodes;
    std::string newick;
    if (N.size() == 1) return "(" + N[0].identifier + ");";
    // nodes are stored in creation order == the pre-order the reference's dfsExpansion yields
    size_t curr_level = 0;
    bool prev_open = true;
    std::vector<const Node*> node_stack;
    auto length = [](float bl) { return bl >= 0 ? ":" + std::to_string(bl) : std::string(); };
    const size_t level_offset = N[0].level - 1;
    for (const Node& n : N) {
        const size_t level = n.level - level_offset;
        if (curr_level < level) {
            if (!prev_open) newick += ',';
            size_t l = level - 1;
            if (curr_level > 1) l = level - curr_level;
            for (size_t i = 0; i < l; i++) {
                newick += '(';
                prev_open = true;
            }
            if (n.children.empty()) {
                newick += n.identifier + length(n.branchLength);
                prev_open = false;
            } else {
                node_stack.push_back(&n);
            }
        } else if (curr_level > level) {
            prev_open = false;
            for (size_t i = level; i < curr_level; i++) {
                newick += ')' + node_stack.back()->identifier + length(node_stack.back()->branchLength);
                node_stack.pop_back();
            }
            if (n.children.empty()) newick += ',' + n.identifier + length(n.branchLength);
            else node_stack.push_back(&n);
        } else {
            prev_open = false;
            if (n.children.empty()) newick += ',' + n.identifier + length(n.branchLength);
            else node_stack.push_back(&n);
        }
        curr_level = level;
    }
    while (!node_stack.empty()) {
        newick += ')' + node_stack.back()->identifier + length(node_stack.back()->branchLength);
        node_stack.pop_back();
    }
    newick += ';';
    return newick;
}

namespace {

constexpr size_t kNewickChunk = 1u << 20;  // NEWICK_CHUNK_SIZE, src/panman.cpp:44

// One Cap'n Proto segment under construction.  Offsets are in 64-bit words.
class Segment {
  public:
    std::vector<uint64_t> w;
    size_t alloc(size_t words) {
        const size_t at = w.size();
        w.resize(at + words, 0);
        return at;
    }
    // struct pointer at word `p` to a struct of (data, ptrs) words at `target`
    void struct_ptr(size_t p, size_t target, uint32_t data, uint32_t ptrs) {
        w[p] = offset_bits(p, target, 0) | ((uint64_t)data << 32) | ((uint64_t)ptrs << 48);
    }
    // list pointer: elem = 1 bit, 2 byte, 4 four bytes, 5 eight bytes, 6 pointer, 7 composite (count = words after the tag)
    void list_ptr(size_t p, size_t target, uint32_t elem, uint64_t count) {
        if (count >= (1ull << 29)) throw BuildError(-20, "panman writer: list too long for one Cap'n Proto list pointer");
        w[p] = offset_bits(p, target, 1) | ((uint64_t)elem << 32) | (count << 35);
    }
    // struct list: tag word + n elements of (data, ptrs) words; returns the word of element 0
    size_t struct_list(size_t p, size_t n, uint32_t data, uint32_t ptrs) {
        const size_t tag = alloc(1 + n * (data + ptrs));
        list_ptr(p, tag, 7, (uint64_t)n * (data + ptrs));
        w[tag] = ((uint64_t)n << 2) | ((uint64_t)data << 32) | ((uint64_t)ptrs << 48);
        return tag + 1;
    }
    void text(size_t p, const char* s, size_t len) {  // List(UInt8) with the terminating NUL counted
        const size_t at = alloc((len + 1 + 7) / 8);
        memcpy((char*)&w[at], s, len);
        list_ptr(p, at, 2, len + 1);
    }
    template <class T>
    T* data_list(size_t p, size_t n, uint32_t elem) {
        const size_t at = alloc((n * sizeof(T) + 7) / 8);
        list_ptr(p, at, elem, n);
        return (T*)&w[at];
    }

  private:
    uint64_t offset_bits(size_t p, size_t target, uint64_t kind) const {
        const int64_t off = (int64_t)target - (int64_t)p - 1;
        if (off < -(1ll << 29) || off >= (1ll << 29))
            throw BuildError(-20, "panman writer: message too large for one Cap'n Proto segment (4 GB)");
        return kind | (((uint64_t)off & 0x3fffffffull) << 2);
    }
};

}  // namespace

// The Cap'n Proto byte stream (segment table + one segment) of TreeGroup{trees = [this tree], complexMutations = []}.
std::vector<uint8_t> encodePanMAN(const Tree& T) {
    Segment S;
    S.alloc(1);                                   // root pointer
    const size_t tg = S.alloc(2);                 // TreeGroup: 0 data words, 2 pointers (trees, complexMutations)
    S.struct_ptr(0, tg, 0, 2);
    const size_t tree = S.struct_list(tg + 0, 1, 0, 8);  // Tree: 8 pointers
    // pointer slots of Tree: 0 newick, 1 nodes, 2 consensusSeqMap, 3 gaps, 4 blockGaps, 5 circularSequences,
    // 6 rotationIndexes, 7 sequencesInverted

    // nodes: allNodes.size() + 1 entries (src/panman.cpp:7039), pre-order; the last one stays empty
    const size_t V = T.nodes.size();
    const size_t nodes = S.struct_list(tree + 1, V + 1, 0, 2);  // Node: mutations, annotations
    for (size_t v = 0; v < V; v++) {
        const Node& n = T.nodes[v];
        const size_t np = nodes + 2 * v;
        // blockToMutations (src/panman.cpp:2871-2899): the MSA path has the single block (0,-1); it appears when the node
        // has nucleotide mutations (marker 2 = no block mutation) and/or a block mutation (the root's BI)
        const bool has_block_mut = n.blockInsertionAtRoot, has_nuc = !n.nucMutation.empty();
        const size_t n_mut = (has_block_mut || has_nuc) ? 1 : 0;
        const size_t m = S.struct_list(np + 0, n_mut, 2, 1);  // Mutation: 2 data words, 1 pointer
        if (n_mut) {
            // data word 0: blockId = primaryBlockId << 32 (secondaryBlockId == -1: blockGapExist = false)
            S.w[m + 0] = 0;
            // data word 1 bits: 0 blockGapExist, 1 blockMutExist, 2 blockMutInfo, 3 blockInversion (:2903-2910):
            // with a block mutation (BI, not inverted): exist, info = true, inversion = false; without: marker 2 ->
            // exist = false, info = bool(2) = true, inversion = true
            uint64_t bits = 0;
            if (has_block_mut) bits |= (1u << 1) | (1u << 2);
            else bits |= (1u << 2) | (1u << 3);
            S.w[m + 1] = bits;
            const size_t k = n.nucMutation.size();
            const size_t nm = S.struct_list(m + 2, k, 2, 0);  // NucMut: 2 data words
            for (size_t i = 0; i < k; i++) {
                const NucMut& x = n.nucMutation[i];
                const uint32_t len = x.mutInfo >> 4;
                const uint32_t info = ((x.nucs >> (24 - len * 4)) << 8) + x.mutInfo;  // :2890
                const bool gap = x.nucGapPosition != -1;
                S.w[nm + 2 * i] = (uint64_t)(uint32_t)x.nucPosition | ((uint64_t)(gap ? (uint32_t)x.nucGapPosition : 0u) << 32);
                S.w[nm + 2 * i + 1] = (gap ? 1ull : 0ull) | ((uint64_t)info << 32);
            }
        }
        S.list_ptr(np + 1, S.alloc(0), 6, 0);  // annotations: empty List(Text)
    }

    // newick in chunks (splitNewick)
    {
        const std::string nw = getNewickString(T);
        const size_t chunks = (nw.size() + kNewickChunk - 1) / kNewickChunk;
        const size_t at = S.alloc(chunks);
        S.list_ptr(tree + 0, at, 6, chunks);
        for (size_t i = 0; i < chunks; i++) {
            const size_t b = i * kNewickChunk, len = std::min(kNewickChunk, nw.size() - b);
            S.text(at + i, nw.data() + b, len);
        }
    }
    // consensusSeqMap: one entry {blockId [0], consensusSeq words, blockGapExist [false]} (:7049-7086)
    {
        const size_t c = S.struct_list(tree + 2, 1, 0, 4);  // ConsensusSeqToBlockIds: 4 pointers
        int64_t* ids = S.data_list<int64_t>(c + 0, 1, 5);
        ids[0] = 0;
        uint32_t* words = S.data_list<uint32_t>(c + 1, T.blockConsensusSeq.size(), 4);
        if (!T.blockConsensusSeq.empty()) memcpy(words, T.blockConsensusSeq.data(), T.blockConsensusSeq.size() * sizeof(uint32_t));
        S.list_ptr(c + 2, S.alloc(1), 1, 1);  // blockGapExist = [false]: a one-element List(Bool) is one word of bits
        // chromosomeName is never set: null pointer
    }
    S.struct_list(tree + 3, 0, 2, 2);  // gaps: none on this path (GapList: 2 data, 2 pointers)
    // blockGaps (slot 4) is not touched by TreeGroup::writeToFile: null
    S.struct_list(tree + 5, 0, 1, 1);  // circularSequences
    S.struct_list(tree + 6, 0, 1, 1);  // rotationIndexes
    S.struct_list(tree + 7, 0, 1, 1);  // sequencesInverted
    S.struct_list(tg + 1, 0, 10, 3);   // complexMutations: none (ComplexMutation: 10 data words, 3 pointers)

    // stream framing: (segment count - 1) u32, segment size in words u32, then the segment
    if (S.w.size() >= (1ull << 32)) throw BuildError(-20, "panman writer: segment too large");
    std::vector<uint8_t> out(8 + S.w.size() * 8);
    const uint32_t head[2] = {0u, (uint32_t)S.w.size()};
    memcpy(out.data(), head, 8);
    memcpy(out.data() + 8, S.w.data(), S.w.size() * 8);
    return out;
}

// `path` gets the .panman file.  compress = true: through xz -9 like the reference's CLI; false: the bare message.
void writePanMAN(const Tree& T, const std::string& path, bool compress) {
    const std::vector<uint8_t> msg = encodePanMAN(T);
    if (!compress) {
        std::ofstream f(path, std::ios::binary);
        if (!f.good()) throw BuildError(-11, "panman writer: cannot open " + path);
        f.write((const char*)msg.data(), (std::streamsize)msg.size());
        return;
    }
    std::string quoted = "'";
    for (char ch : path) quoted += ch == '\'' ? std::string("'\\''") : std::string(1, ch);
    quoted += "'";
    FILE* p = popen(("xz -9 -c > " + quoted).c_str(), "w");
    if (!p) throw BuildError(-11, "panman writer: cannot start xz");
    const size_t n = fwrite(msg.data(), 1, msg.size(), p);
    const int rc = pclose(p);
    if (n != msg.size() || rc != 0) throw BuildError(-11, "panman writer: xz failed for " + path);
}

}  // namespace pmfs_host

// ---- C ABI (include/pmfs_host.h) ---------------------------------------------------------------------------------------
namespace {
thread_local std::string g_writer_error;
}

extern "C" {

const char* pmfs_host_writer_last_error(void) { return g_writer_error.c_str(); }

// The writer alone, from arrays (no device needed): a tree in pre-order (parent[], '\n'-joined identifiers, branch
// lengths or NULL for 0.0), the per-node packed NucMut lists as a CSR, the Block consensus words.
int pmfs_host_write_panman(int32_t n_nodes, const int32_t* parent, const char* names_joined, const float* branch_length,
                           const int64_t* nuc_off, const pmfs_host_nucmut_t* nuc, const uint32_t* block_words,
                           int64_t n_block_words, const char* out_path, int compress) {
    using namespace pmfs_host;
    try {
        if (n_nodes < 1 || !parent || !names_joined || !nuc_off || !out_path) throw BuildError(-1, "pmfs_host_write_panman: bad argument");
        // (Tree has no default constructor on purpose -- it IS the reference's constructor; the writer only reads these fields)
        struct Bare : Tree {
            Bare() : Tree(nullptr) {}
        } T;
        T.nodes.resize((size_t)n_nodes);
        const char* s = names_joined;
        for (int32_t v = 0; v < n_nodes; v++) {
            const char* e = strchr(s, '\n');
            if (!e) throw BuildError(-1, "pmfs_host_write_panman: fewer names than nodes");
            Node& n = T.nodes[(size_t)v];
            n.identifier.assign(s, e);
            s = e + 1;
            n.parent = parent[v];
            if (v > 0 && (parent[v] < 0 || parent[v] >= v)) throw BuildError(-1, "pmfs_host_write_panman: parent[i] must be in [0, i)");
            n.level = v == 0 ? 1 : T.nodes[(size_t)parent[v]].level + 1;
            n.branchLength = branch_length ? branch_length[v] : 0.0f;
            if (v > 0) T.nodes[(size_t)parent[v]].children.push_back(v);
            for (int64_t k = nuc_off[v]; k < nuc_off[v + 1]; k++) {
                NucMut m;
                m.nucPosition = nuc[k].nucPosition; m.nucGapPosition = nuc[k].nucGapPosition;
                m.primaryBlockId = nuc[k].primaryBlockId; m.secondaryBlockId = nuc[k].secondaryBlockId;
                m.mutInfo = (uint8_t)nuc[k].mutInfo; m.nucs = nuc[k].nucs;
                n.nucMutation.push_back(m);
            }
        }
        T.nodes[0].blockInsertionAtRoot = true;  // src/panman.cpp:1430
        T.blockConsensusSeq.assign(block_words, block_words + n_block_words);
        writePanMAN(T, out_path, compress != 0);
        return 0;
    } catch (const pmfs_host::BuildError& e) {
        g_writer_error = e.what();
        return e.status;
    } catch (const std::exception& e) {
        g_writer_error = e.what();
        return -19;
    }
}

}  // extern "C"
