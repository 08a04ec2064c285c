#include "msa_tree.hpp"

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <queue>
#include <stack>

namespace pmfs_host {

char getNucleotideFromCode(int code) {  // src/panman.cpp:46-81
    static const char tab[16] = {'-', 'A', 'C', 'M', 'G', 'R', 'S', 'V', 'T', 'W', 'Y', 'H', 'K', 'D', 'B', 'N'};
    return (code >= 1 && code <= 15) ? tab[code] : '-';
}

char getCodeFromNucleotide(char nuc) {  // src/panman.cpp:83-118 (case-sensitive: lower case is a gap)
    switch (nuc) {
        case 'A': return 1;  case 'C': return 2;  case 'M': return 3;  case 'G': return 4;
        case 'R': return 5;  case 'S': return 6;  case 'V': return 7;  case 'T': return 8;
        case 'W': return 9;  case 'Y': return 10; case 'H': return 11; case 'K': return 12;
        case 'D': return 13; case 'B': return 14; case 'N': return 15;
        default: return 0;
    }
}

// Quote-aware split (src/panman.cpp:255-285): a piece holding an odd number of apostrophes is glued to the following
// pieces until the count is even again; an empty trailing piece is dropped.
void stringSplit(std::string const& s, char delim, std::vector<std::string>& words) {
    size_t from = 0, hit, glueFrom = 0;
    bool glue = false;
    auto oddQuotes = [](const std::string& t) { return std::count(t.begin(), t.end(), '\'') % 2 == 1; };
    while ((hit = s.find(delim, from)) != std::string::npos) {
        if (!glue) {
            std::string piece = s.substr(from, hit - from);
            // the reference tests its glue marker against 0, so a glued run that would start at offset 0 is never
            // detected as one: the piece is then silently dropped together with its successors up to the next
            // even-quoted piece.  Reproduced: glueFrom == 0 means "not gluing".
            if (oddQuotes(piece)) {
                glueFrom = from;
                glue = glueFrom != 0;
            } else {
                words.emplace_back(std::move(piece));
            }
        } else {
            std::string piece = s.substr(glueFrom, hit - glueFrom);
            if (!oddQuotes(piece)) {
                glue = false;
                glueFrom = 0;
                words.emplace_back(std::move(piece));
            }
        }
        from = hit + 1;
    }
    std::string last = s.substr(from);
    if (!last.empty()) words.push_back(std::move(last));
}

std::string stripString(std::string s) {  // src/panman.cpp:288-298
    size_t e = s.size();
    while (e > 0 && s[e - 1] == ' ') e--;
    s.resize(e);
    size_t b = 0;
    while (b < s.size() && s[b] == ' ') b++;
    return b < s.size() ? s.substr(b) : s;
}

// src/panman.cpp:300-440.  Internal labels are ignored; internal nodes are named node_1, node_2, ... in the order
// their '(' appears (src/panman.hpp:426-428); a missing or zero branch length becomes 1.0; the root's is 0.
void Tree::createTreeFromNewickString(std::string newick, std::vector<Node>& nodes,
                                      std::unordered_map<std::string, int32_t>& allNodes) {
    newick = stripString(newick);
    struct Token {
        std::string leaf;
        size_t opens = 0, closes = 0;
    };
    std::vector<Token> tokens;
    std::vector<std::queue<float>> branchLen(128);
    size_t level = 0;
    std::vector<std::string> pieces;
    stringSplit(newick, ',', pieces);
    auto lengthOf = [](const std::string& digits) {
        float len = digits.empty() ? 1.0f : std::stof(digits);
        return len == 0 ? 1.0f : len;
    };
    for (const std::string& piece : pieces) {
        Token tk;
        bool pastName = false, inLength = false, quoted = false, sawQuote = false;
        std::string digits;
        for (char ch : piece) {
            if (quoted) {
                tk.leaf += ch;
                if (ch == '\'') quoted = false;
            } else if (ch == '\'') {
                quoted = sawQuote = true;
                tk.leaf += ch;
            } else if (ch == ':') {
                pastName = true;
                digits.clear();
                inLength = true;
            } else if (ch == '(') {
                tk.opens++;
                level++;
                if (branchLen.size() <= level) branchLen.resize(level * 2);
            } else if (ch == ')') {
                pastName = true;
                tk.closes++;
                if (level == 0) throw BuildError(-10, "incorrect Newick format: unbalanced ')'");
                branchLen[level].push(lengthOf(digits));
                level--;
                inLength = false;
            } else if (!pastName) {
                tk.leaf += ch;
                inLength = false;
            } else if (inLength && (isdigit((unsigned char)ch) || ch == '.')) {
                digits += ch;
            }
        }
        if (sawQuote && tk.leaf.size() >= 2 && tk.leaf.front() == '\'' && tk.leaf.back() == '\'')
            tk.leaf = tk.leaf.substr(1, tk.leaf.size() - 2);
        branchLen[level].push(lengthOf(digits));
        tokens.push_back(std::move(tk));
    }
    if (level != 0) throw BuildError(-10, "incorrect Newick format: unbalanced '('");

    nodes.clear();
    allNodes.clear();
    std::vector<int32_t> open;  // stack of internal nodes still accepting children
    size_t internalCount = 0;
    auto add = [&](const std::string& id, float len) {
        Node n;
        n.identifier = id;
        n.branchLength = len;
        n.parent = open.empty() ? -1 : open.back();
        n.level = n.parent < 0 ? 1 : nodes[n.parent].level + 1;
        int32_t idx = (int32_t)nodes.size();
        if (n.parent >= 0) nodes[n.parent].children.push_back(idx);
        allNodes[id] = idx;
        nodes.push_back(std::move(n));
        return idx;
    };
    for (const Token& tk : tokens) {
        for (size_t j = 0; j < tk.opens; j++) {
            int32_t idx = add("node_" + std::to_string(++internalCount), branchLen[level].front());
            branchLen[level].pop();
            level++;
            open.push_back(idx);
        }
        if (open.empty()) throw BuildError(-10, "Newick tree has no internal node");
        add(tk.leaf, branchLen[level].front());
        branchLen[level].pop();
        for (size_t j = 0; j < tk.closes; j++) {
            open.pop_back();
            level--;
        }
    }
    if (nodes.empty()) throw BuildError(-10, "tree found empty");
    nodes[0].branchLength = 0.0;
}

// Default mode: src/panman.cpp:1277-1315 ('\r' stripped from rows and from all ids but the last one).
// Low-mem mode: the length scan of :1469-1490 plus readFastaInBatch :667-714 (no '\r' handling at all).
void Tree::readMSA(std::istream& fin, FILE_TYPE ftype, std::map<std::string, std::string>& seqs, size_t& lineLength) {
    const bool strip = ftype == FILE_TYPE::MSA;
    auto firstPiece = [](const std::string& s, char delim) {
        std::vector<std::string> w;
        stringSplit(s, delim, w);
        return w.empty() ? std::string() : w[0];
    };
    std::string line, seq, id;
    lineLength = 0;
    auto mismatch = [&]() { throw BuildError(-11, "sequence lengths don't match: " + id); };
    while (getline(fin, line, '\n')) {
        if (line.empty()) continue;
        if (line[0] == '>') {
            if (!seq.empty()) {
                if (lineLength == 0) lineLength = seq.size();
                else if (lineLength != seq.size()) mismatch();
                seqs[strip ? firstPiece(id, '\r') : id] = seq;
            }
            id = firstPiece(line, ' ').substr(1);
            seq.clear();
        } else {
            seq += strip ? firstPiece(line, '\r') : line;
        }
    }
    if (!seq.empty()) {
        if (lineLength != 0 && lineLength != seq.size()) mismatch();
        lineLength = seq.size();
        seqs[id] = seq;
    }
}

// Fitch branch (src/panman.cpp:1322-1352): the --reference row, else the first non-gap character in map (id-sorted)
// order.  Sankoff branch (:1517-1547): the reference character unless it is a gap, then the first non-gap.
std::string Tree::computeConsensus(const std::map<std::string, std::string>& seqs, size_t lineLength, FILE_TYPE ftype,
                                   const std::string& reference) {
    auto firstNonGap = [&](size_t i, char& out) {
        for (const auto& u : seqs)
            if (u.second[i] != '-') {
                out = u.second[i];
                return true;
            }
        return false;
    };
    std::string cons;
    const auto ref = reference.empty() ? seqs.end() : seqs.find(reference);
    if (ftype == FILE_TYPE::MSA) {
        if (!reference.empty()) return ref == seqs.end() ? std::string() : ref->second;  // absent id: zero columns, as upstream
        cons.resize(lineLength);
        for (size_t i = 0; i < lineLength; i++)
            if (!firstNonGap(i, cons[i]))
                throw BuildError(-12, "column " + std::to_string(i) + " holds only gaps (undefined behaviour in the reference)");
        return cons;
    }
    cons.assign(lineLength, '\0');
    if (!reference.empty()) {
        if (ref == seqs.end()) throw BuildError(-13, "Reference not found in the sequence");
        for (size_t i = 0; i < lineLength; i++) {
            if (ref->second[i] == '-') firstNonGap(i, cons[i]);
            else cons[i] = ref->second[i];
        }
    } else {
        for (size_t i = 0; i < lineLength; i++)
            if (!firstNonGap(i, cons[i])) throw BuildError(-12, "column " + std::to_string(i) + " holds only gaps");
    }
    return cons;
}

NucMut::NucMut(const pmfs_mut* a, int start, int end) {
    primaryBlockId = 0;
    secondaryBlockId = -1;
    mutInfo = (uint8_t)(((end - start) << 4) + a[start].type);
    nucPosition = a[start].pos;
    nucGapPosition = -1;
    nucs = 0;
    for (int i = start; i < end; i++) nucs += ((uint32_t)a[i].code << (4 * (5 - (i - start))));
}

// src/panman.cpp:1435-1456: a new record starts when the run reaches 6, the position is not the successor, or the
// type changes.  Records arrive sorted by (node,pos), which is the order std::sort gives each node's tuple list.
void Tree::packNucMutations(const std::vector<pmfs_mut>& m, std::vector<Node>& nodes) {
    size_t a = 0;
    while (a < m.size()) {
        size_t b = a;
        while (b < m.size() && m[b].node == m[a].node) b++;
        std::vector<NucMut>& out = nodes[m[a].node].nucMutation;
        size_t start = a;
        for (size_t i = a + 1; i < b; i++)
            if (i - start == 6 || m[i].pos != m[i - 1].pos + 1 || m[i].type != m[i - 1].type) {
                out.emplace_back(m.data(), (int)start, (int)i);
                start = i;
            }
        out.emplace_back(m.data(), (int)start, (int)b);
        a = b;
    }
}

std::vector<uint32_t> Tree::packBlockConsensus(const std::string& seq) {  // src/panman.cpp:236-247
    std::vector<uint32_t> words;
    for (size_t i = 0; i < seq.size(); i += 8) {
        uint32_t w = 0;
        for (size_t j = i; j < std::min(i + 8, seq.size()); j++) w ^= ((uint32_t)getCodeFromNucleotide(seq[j]) << (4 * (7 - (j - i))));
        words.push_back(w);
    }
    return words;
}

Tree::Tree(std::istream& fin, std::istream& secondFin, FILE_TYPE ftype, std::string reference, int device) {
    if (ftype != FILE_TYPE::MSA && ftype != FILE_TYPE::MSA_OPTIMIZE)
        throw BuildError(PMFS_ERR_ARG, "only FILE_TYPE::MSA and MSA_OPTIMIZE are built by this library");
    std::string newickString;
    std::getline(secondFin, newickString);
    createTreeFromNewickString(newickString, nodes, allNodes);

    std::map<std::string, std::string> seqs;
    size_t lineLength = 0;
    readMSA(fin, ftype, seqs, lineLength);
    consensusSeq = computeConsensus(seqs, lineLength, ftype, reference);
    const int64_t L = (int64_t)consensusSeq.size();

    // flat problem: rows in map order, leaves looked up by name (the reference keys its per-column maps by name)
    const int32_t V = (int32_t)nodes.size();
    std::unordered_map<std::string, int32_t> rowOf;
    int32_t r = 0;
    for (const auto& u : seqs) rowOf[u.first] = r++;
    std::vector<int32_t> parent(V), childOff(V + 1, 0), childIdx, leafRow(V);
    for (int32_t v = 0; v < V; v++) {
        parent[v] = nodes[v].parent;
        childOff[v + 1] = childOff[v] + (int32_t)nodes[v].children.size();
        childIdx.insert(childIdx.end(), nodes[v].children.begin(), nodes[v].children.end());
        if (!nodes[v].children.empty()) leafRow[v] = -1;
        else {
            auto it = rowOf.find(nodes[v].identifier);
            leafRow[v] = it == rowOf.end() ? -2 : it->second;
        }
    }
    if (L > 0) {
        // the map's strings as they are, in map order: staged, uploaded and converted to packed codes chunk by chunk
        // (pmfs_run_char_rows; the character -> code conversion of :1398-1406 / :1565-1571 happens on the device)
        std::vector<const char*> rowPtrs;
        rowPtrs.reserve(seqs.size());
        for (const auto& u : seqs) rowPtrs.push_back(u.second.data());
        const char* forcedChars = nullptr;
        if (!reference.empty()) forcedChars = seqs.at(reference).data();  // refState (:1409) / defaultState (:1573-1580)
        pmfs_tree t = {V, parent.data(), childOff.data(), childIdx.data(), leafRow.data()};
        pmfs_ctx* ctx = nullptr;
        int rc = pmfs_create(&ctx, &t, device);
        if (rc) throw BuildError(rc, pmfs_last_error());
        pmfs_mut* out = nullptr;
        int64_t n = 0;
        rc = pmfs_run_char_rows(ctx, ftype == FILE_TYPE::MSA ? PMFS_FITCH : PMFS_SANKOFF, rowPtrs.data(), (int32_t)seqs.size(), L,
                                consensusSeq.data(), forcedChars, &out, &n);
        if (rc) {
            std::string msg = pmfs_last_error();
            pmfs_destroy(ctx);
            throw BuildError(rc, msg);
        }
        mutations.assign(out, out + n);
        pmfs_get_timing(ctx, &timing);
        pmfs_result_free(ctx, out);
        pmfs_destroy(ctx);
    }
    blockConsensusSeq = packBlockConsensus(consensusSeq);  // blocks.emplace_back(0, consensusSeq)
    nodes[0].blockInsertionAtRoot = true;
    packNucMutations(mutations, nodes);
}

}  // namespace pmfs_host

// ---- C ABI over the host mirror (tests, and the `pmbuild`-style callers that are not C++) ------------------------------
using pmfs_host::BuildError;
using pmfs_host::Tree;

struct pmfs_host_tree {
    Tree* tree;
};

static thread_local std::string g_host_error;

extern "C" {

const char* pmfs_host_last_error(void) { return g_host_error.c_str(); }

// Newick only (no device needed): returns node count or a negative status.  Arrays may be null to query sizes.
int pmfs_host_parse_newick(const char* newick, int32_t* parent, int32_t* n_children, char* names, int64_t names_cap) {
    try {
        std::vector<pmfs_host::Node> nodes;
        std::unordered_map<std::string, int32_t> all;
        Tree::createTreeFromNewickString(newick, nodes, all);
        std::string joined;
        for (auto& n : nodes) joined += n.identifier + "\n";
        if (parent)
            for (size_t i = 0; i < nodes.size(); i++) parent[i] = nodes[i].parent;
        if (n_children)
            for (size_t i = 0; i < nodes.size(); i++) n_children[i] = (int32_t)nodes[i].children.size();
        if (names) {
            if ((int64_t)joined.size() + 1 > names_cap) return -14;
            memcpy(names, joined.c_str(), joined.size() + 1);
        }
        return (int)nodes.size();
    } catch (const BuildError& e) {
        g_host_error = e.what();
        return e.status;
    } catch (const std::exception& e) {
        g_host_error = e.what();
        return -19;
    }
}

void pmfs_host_free(void* p) { free(p); }

// `panmanUtils -M msa -N nwk [--reference id] [--low-mem-mode] -o out` (src/panmanUtils.cpp:1487-1543): build, then
// writePanMAN.
int pmfs_host_build_panman(const char* msa_path, const char* nwk_path, const char* reference, int low_mem, int device,
                           const char* out_path, int compress) {
    try {
        std::ifstream fin(msa_path), nwk(nwk_path);
        if (!fin.good() || !nwk.good() || !out_path) {
            g_host_error = "cannot open input file";
            return -15;
        }
        Tree T(fin, nwk, low_mem ? pmfs_host::MSA_OPTIMIZE : pmfs_host::MSA, reference ? reference : "", device);
        pmfs_host::writePanMAN(T, out_path, compress != 0);
        return 0;
    } catch (const BuildError& e) {
        g_host_error = e.what();
        return e.status;
    } catch (const std::exception& e) {
        g_host_error = e.what();
        return -19;
    }
}

typedef struct {  // NucMut (src/panman.hpp:55-61) with mutInfo widened so the layout has no padding
    int32_t nucPosition, nucGapPosition, primaryBlockId, secondaryBlockId;
    uint32_t mutInfo, nucs;
} pmfs_host_nucmut;

// The whole `panmanUtils -M msa -N nwk [--reference id] [--low-mem-mode]` construction up to (not including)
// serialisation, on CUDA device `device`.  Outputs are malloc'd (release with pmfs_host_free):
//   names: '\n'-joined node identifiers in pre-order; parent[V]; consensus characters; records sorted by (node,pos);
//   per-node NucMut lists as CSR (nuc_off[V+1], nuc[]); Block(0, consensus) words.
int pmfs_host_build(const char* msa_path, const char* nwk_path, const char* reference, int low_mem, int device,
                    int32_t* n_nodes, int32_t** parent, char** names, int64_t* n_cols, char** consensus, pmfs_mut** muts,
                    int64_t* n_muts, int64_t** nuc_off, pmfs_host_nucmut** nuc, int64_t* n_nuc, uint32_t** block_words,
                    int64_t* n_block_words) {
    try {
        std::ifstream fin(msa_path), nwk(nwk_path);
        if (!fin.good() || !nwk.good()) {
            g_host_error = "cannot open input file";
            return -15;
        }
        Tree T(fin, nwk, low_mem ? pmfs_host::MSA_OPTIMIZE : pmfs_host::MSA, reference ? reference : "", device);
        const size_t V = T.nodes.size();
        auto dup = [](const void* src, size_t bytes) {
            void* p = malloc(bytes ? bytes : 1);
            if (bytes) memcpy(p, src, bytes);
            return p;
        };
        std::vector<int32_t> par(V);
        std::string joined;
        std::vector<int64_t> off(V + 1, 0);
        std::vector<pmfs_host_nucmut> packed;
        for (size_t v = 0; v < V; v++) {
            par[v] = T.nodes[v].parent;
            joined += T.nodes[v].identifier + "\n";
            off[v] = (int64_t)packed.size();
            for (const auto& m : T.nodes[v].nucMutation)
                packed.push_back({m.nucPosition, m.nucGapPosition, m.primaryBlockId, m.secondaryBlockId, m.mutInfo, m.nucs});
        }
        off[V] = (int64_t)packed.size();
        *n_nodes = (int32_t)V;
        *parent = (int32_t*)dup(par.data(), V * sizeof(int32_t));
        *names = (char*)dup(joined.c_str(), joined.size() + 1);
        *n_cols = (int64_t)T.consensusSeq.size();
        *consensus = (char*)dup(T.consensusSeq.c_str(), T.consensusSeq.size() + 1);
        *muts = (pmfs_mut*)dup(T.mutations.data(), T.mutations.size() * sizeof(pmfs_mut));
        *n_muts = (int64_t)T.mutations.size();
        *nuc_off = (int64_t*)dup(off.data(), off.size() * sizeof(int64_t));
        *nuc = (pmfs_host_nucmut*)dup(packed.data(), packed.size() * sizeof(pmfs_host_nucmut));
        *n_nuc = (int64_t)packed.size();
        *block_words = (uint32_t*)dup(T.blockConsensusSeq.data(), T.blockConsensusSeq.size() * sizeof(uint32_t));
        *n_block_words = (int64_t)T.blockConsensusSeq.size();
        return 0;
    } catch (const BuildError& e) {
        g_host_error = e.what();
        return e.status;
    } catch (const std::exception& e) {
        g_host_error = e.what();
        return -19;
    }
}

// Packing only: records sorted by (node,pos) -> per-node NucMut CSR (no device needed).
int pmfs_host_pack(const pmfs_mut* muts, int64_t n_muts, int32_t n_nodes, int64_t** nuc_off, pmfs_host_nucmut** nuc,
                   int64_t* n_nuc) {
    std::vector<pmfs_host::Node> nodes((size_t)n_nodes);
    std::vector<pmfs_mut> v(muts, muts + n_muts);
    for (const auto& m : v)
        if (m.node < 0 || m.node >= n_nodes) return PMFS_ERR_ARG;
    Tree::packNucMutations(v, nodes);
    std::vector<int64_t> off((size_t)n_nodes + 1, 0);
    std::vector<pmfs_host_nucmut> packed;
    for (int32_t i = 0; i < n_nodes; i++) {
        off[i] = (int64_t)packed.size();
        for (const auto& m : nodes[i].nucMutation)
            packed.push_back({m.nucPosition, m.nucGapPosition, m.primaryBlockId, m.secondaryBlockId, m.mutInfo, m.nucs});
    }
    off[n_nodes] = (int64_t)packed.size();
    *nuc_off = (int64_t*)malloc(off.size() * sizeof(int64_t));
    memcpy(*nuc_off, off.data(), off.size() * sizeof(int64_t));
    *nuc = (pmfs_host_nucmut*)malloc(std::max<size_t>(1, packed.size()) * sizeof(pmfs_host_nucmut));
    if (!packed.empty()) memcpy(*nuc, packed.data(), packed.size() * sizeof(pmfs_host_nucmut));
    *n_nuc = (int64_t)packed.size();
    return 0;
}
}
