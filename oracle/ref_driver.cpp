// TEST INFRASTRUCTURE ONLY (oracle/_ref). Not part of the product: only tests/, __graft_entry__.smoke() and
// bench.py's cpu_baseline / --impl reference legs may load the library built from this file.
//
// What this is: a driver around the REFERENCE'S OWN Fitch / Sankoff code.  oracle/Makefile compiles
// /root/reference/src/fitchSankoff.cpp unmodified (in place, via stdin) against oracle/shim/panmanUtils.hpp and
// links it with this file into oracle/_ref/libpmref.so.  All per-column arithmetic therefore IS the reference.
// The reference's construction driver (src/panman.cpp) cannot be compiled here (TBB/Boost/capnp/jsoncpp absent),
// so the parts of it that surround the three per-column calls are restated below, each citing the lines it follows:
//   * code tables                      src/panman.cpp:46-81, :83-118
//   * stringSplit / stripString        src/panman.cpp:255-298
//   * createTreeFromNewickString       src/panman.cpp:300-440   (node naming: src/panman.hpp:426-428)
//   * MSA reader (default mode)        src/panman.cpp:1271-1315
//   * MSA reader (low-mem mode)        src/panman.cpp:667-714, :1469-1490
//   * consensus                        src/panman.cpp:1322-1352 (Fitch branch), :1517-1547 (Sankoff branch)
//   * per-column loop, Fitch           src/panman.cpp:1371-1425
//   * per-column loop, Sankoff         src/panman.cpp:1558-1603
//   * sort + <=6-run packing           src/panman.cpp:1435-1456, :1615-1636; NucMut range ctor src/panman.hpp:86-128
//   * Block consensus packing          src/panman.cpp:236-247
// Differences from the reference that are deliberate and documented in DESIGN.md:
//   * exit()/assert() paths are turned into negative return codes (a test process must survive them);
//   * columns are processed by std::thread workers with large stacks (TBB is absent; the reference recursion
//     overflows an 8 MB stack on a 100 000-deep ladder);
//   * an all-gap column without --reference is rejected (the reference's behaviour there is undefined, SURVEY §8 notes).
#include "shim/panmanUtils.hpp"

#include <pthread.h>

#include <atomic>
#include <cctype>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <mutex>
#include <queue>
#include <set>
#include <stack>
#include <tuple>

using panmanUtils::Node;
using panmanUtils::Tree;

// ---------------------------------------------------------------- code tables (src/panman.cpp:46-118)
char panmanUtils::getNucleotideFromCode(int code) {
    static const char tab[16] = {'-', 'A', 'C', 'M', 'G', 'R', 'S', 'V', 'T', 'W', 'Y', 'H', 'K', 'D', 'B', 'N'};
    return (code >= 1 && code <= 15) ? tab[code] : '-';
}
char panmanUtils::getCodeFromNucleotide(char nuc) {
    switch (nuc) {
        case 'A': return 1;  case 'C': return 2;  case 'G': return 4;  case 'T': return 8;
        case 'R': return 5;  case 'Y': return 10; case 'S': return 6;  case 'W': return 9;
        case 'K': return 12; case 'M': return 3;  case 'B': return 14; case 'D': return 13;
        case 'H': return 11; case 'V': return 7;  case 'N': return 15;
        default: return 0;
    }
}

namespace {

struct RefMut {  // one (node, column) annotation, the tuple of src/panman.cpp:1418 plus the node it belongs to
    int32_t node;
    int32_t pos;
    uint8_t type;  // NS=0 ND=1 NI=2
    uint8_t code;  // getCodeFromNucleotide(mutation char)
    uint8_t pad[2];
};

struct RefNucMut {  // src/panman.hpp:55-61 field for field
    int32_t nucPosition;
    int32_t nucGapPosition;
    int32_t primaryBlockId;
    int32_t secondaryBlockId;
    uint32_t mutInfo;  // uint8_t in the reference; widened for a padding-free C layout
    uint32_t nucs;
};

struct NodeGraph {
    std::vector<Node*> nodes;  // creation order == pre-order; nodes[0] is the root
    std::unordered_map<std::string, int> index;
    ~NodeGraph() {
        for (auto* n : nodes) delete n;
    }
};

// ---------------------------------------------------------------- stringSplit (src/panman.cpp:255-285)
void stringSplitRef(const std::string& s, char delim, std::vector<std::string>& words) {
    size_t start_pos = 0, end_pos = 0, temp_pos = 0;
    while ((end_pos = s.find(delim, start_pos)) != std::string::npos) {
        if (end_pos >= s.length()) break;
        std::string sub;
        if (temp_pos == 0) {
            sub = s.substr(start_pos, end_pos - start_pos);
            if (std::count(sub.begin(), sub.end(), '\'') % 2 == 1) temp_pos = start_pos;
            else words.emplace_back(sub);
        } else {
            sub = s.substr(temp_pos, end_pos - temp_pos);
            if (std::count(sub.begin(), sub.end(), '\'') % 2 == 0) {
                temp_pos = 0;
                words.emplace_back(sub);
            }
        }
        start_pos = end_pos + 1;
    }
    auto last = s.substr(start_pos, s.size() - start_pos);
    if (last != "") words.push_back(std::move(last));
}

std::string stripStringRef(std::string s) {  // src/panman.cpp:288-298
    while (s.length() && s[s.length() - 1] == ' ') s.pop_back();
    for (size_t i = 0; i < s.length(); i++)
        if (s[i] != ' ') return s.substr(i);
    return s;
}

// ---------------------------------------------------------------- Newick (src/panman.cpp:300-440)
// Returns 0, or -1 (unbalanced parentheses: the reference exit(1)s), -2 (empty tree: the reference dereferences null).
int newickToGraph(std::string newick, NodeGraph& g) {
    newick = stripStringRef(newick);
    std::vector<std::string> leaves;
    std::vector<size_t> numOpen, numClose;
    std::vector<std::queue<float>> branchLen(128);
    size_t level = 0;
    std::vector<std::string> s1;
    stringSplitRef(newick, ',', s1);
    for (auto s : s1) {
        size_t no = 0, nc = 0;
        bool stop = false, branchStart = false, nameZone = false, hasApo = false;
        std::string leaf, branch;
        for (auto c : s) {
            if (nameZone) {
                leaf += c;
                if (c == '\'') nameZone = false;
            } else if (c == '\'' && !nameZone) {
                nameZone = true;
                hasApo = true;
                leaf += c;
            } else if (c == ':') {
                stop = true;
                branch = "";
                branchStart = true;
            } else if (c == '(') {
                no++;
                level++;
                if (branchLen.size() <= level) branchLen.resize(level * 2);
            } else if (c == ')') {
                stop = true;
                nc++;
                float len = (branch.size() > 0) ? std::stof(branch) : 1.0f;
                if (len == 0) len = 1.0f;
                branchLen[level].push(len);
                if (level == 0) return -1;
                level--;
                branchStart = false;
            } else if (!stop) {
                leaf += c;
                branchStart = false;
            } else if (branchStart) {
                if (isdigit((unsigned char)c) || c == '.') branch += c;
            }
        }
        if (hasApo && leaf[0] == '\'' && leaf[leaf.length() - 1] == '\'') leaf = leaf.substr(1, leaf.length() - 2);
        leaves.push_back(std::move(leaf));
        numOpen.push_back(no);
        numClose.push_back(nc);
        float len = (branch.size() > 0) ? std::stof(branch) : 1.0f;
        if (len == 0) len = 1.0f;
        branchLen[level].push(len);
    }
    if (level != 0) return -1;

    std::stack<Node*> parentStack;
    size_t currInternal = 0;
    auto addNode = [&](const std::string& id, Node* par, float len) {
        Node* n = new Node();
        n->identifier = id;
        n->branchLength = len;
        n->parent = par;
        n->level = par ? par->level + 1 : 1;
        if (par) par->children.push_back(n);
        g.index[id] = (int)g.nodes.size();  // allNodes[id] = node: later duplicates shadow earlier ones, as in the reference
        g.nodes.push_back(n);
        return n;
    };
    for (size_t i = 0; i < leaves.size(); i++) {
        for (size_t j = 0; j < numOpen[i]; j++) {
            std::string nid = "node_" + std::to_string(++currInternal);  // newInternalNodeId, src/panman.hpp:426-428
            Node* nn = addNode(nid, parentStack.empty() ? nullptr : parentStack.top(), branchLen[level].front());
            branchLen[level].pop();
            level++;
            parentStack.push(nn);
        }
        if (parentStack.empty()) return -2;
        addNode(leaves[i], parentStack.top(), branchLen[level].front());
        branchLen[level].pop();
        for (size_t j = 0; j < numClose[i]; j++) {
            parentStack.pop();
            level--;
        }
    }
    if (g.nodes.empty()) return -2;
    g.nodes[0]->branchLength = 0.0;
    return 0;
}

// ---------------------------------------------------------------- MSA readers
// default mode, src/panman.cpp:1277-1315.  Returns 0 or -3 (unequal lengths: the reference exit(-1)s).
int readMsaDefault(std::istream& fin, std::map<std::string, std::string>& seqs, size_t& lineLength) {
    std::string line, currentSequence, currentSequenceId;
    lineLength = 0;
    while (getline(fin, line, '\n')) {
        if (line.length() == 0) continue;
        if (line[0] == '>') {
            if (currentSequence.length()) {
                if (lineLength == 0) lineLength = currentSequence.length();
                else if (lineLength != currentSequence.length()) return -3;
                std::vector<std::string> sp;
                stringSplitRef(currentSequenceId, '\r', sp);
                seqs[sp[0]] = currentSequence;
            }
            std::vector<std::string> sp;
            stringSplitRef(line, ' ', sp);
            currentSequenceId = sp[0].substr(1);
            currentSequence = "";
        } else {
            std::vector<std::string> sp;
            stringSplitRef(line, '\r', sp);
            currentSequence += sp[0];
        }
    }
    if (currentSequence.length()) {
        if (lineLength != 0 && lineLength != currentSequence.length()) return -3;
        lineLength = currentSequence.length();
        seqs[currentSequenceId] = currentSequence;  // NB: the last id is NOT '\r'-stripped in the reference (:1313)
    }
    return 0;
}

// low-mem mode: length scan src/panman.cpp:1469-1490 + readFastaInBatch :667-714 with startIndex=0 and an unbounded
// batch (batching changes I/O only: every column sees the same rows).  No '\r' stripping in this mode.
int readMsaLowMem(std::istream& fin, std::map<std::string, std::string>& seqs, size_t& lineLength) {
    std::string line, currentSequence, currentSequenceId;
    lineLength = 0;
    while (getline(fin, line, '\n')) {
        if (line.length() == 0) continue;
        if (line[0] == '>') {
            if (currentSequence.length()) {
                if (lineLength == 0) lineLength = currentSequence.length();
                else if (lineLength != currentSequence.length()) return -3;
                seqs[currentSequenceId] = currentSequence;
            }
            std::vector<std::string> sp;
            stringSplitRef(line, ' ', sp);
            currentSequenceId = sp[0].substr(1);
            currentSequence = "";
        } else {
            currentSequence += line;
        }
    }
    if (currentSequence.length()) {
        if (lineLength != 0 && lineLength != currentSequence.length()) return -3;
        lineLength = currentSequence.length();
        seqs[currentSequenceId] = currentSequence;
    }
    return 0;
}

// ---------------------------------------------------------------- per-column loops on the reference's members
struct ColumnJob {
    int algo = 0;  // 0 Fitch (src/panman.cpp:1371-1425), 1 Sankoff (:1558-1603)
    Tree* tree = nullptr;
    const NodeGraph* graph = nullptr;
    const std::vector<std::pair<std::string, const char*>>* rows = nullptr;  // (id, row chars), map order
    const char* consensus = nullptr;
    const char* forced = nullptr;  // reference row chars or nullptr
    int64_t nCols = 0;
    std::atomic<int64_t> next{0};
    int64_t chunk = 16;  // columns a thread takes at a time (smaller for small samples, so that every thread gets work)
    std::mutex outMutex;
    std::vector<RefMut>* out = nullptr;
};

void runColumns(ColumnJob* job) {
    using panmanUtils::getCodeFromNucleotide;
    std::vector<RefMut> local;
    Tree& T = *job->tree;
    for (;;) {
        int64_t begin = job->next.fetch_add(job->chunk);
        if (begin >= job->nCols) break;
        int64_t end = std::min(begin + job->chunk, job->nCols);
        for (int64_t i = begin; i < end; i++) {
            std::unordered_map<std::string, std::pair<panmanUtils::NucMutationType, char>> mutations;
            if (job->algo == 0) {
                std::unordered_map<std::string, int> states;
                for (const auto& u : *job->rows) {
                    if (u.second[i] != '-') states.insert({u.first, (1 << getCodeFromNucleotide(u.second[i]))});
                    else states.insert({u.first, 1});
                }
                int refState = job->forced ? (1 << getCodeFromNucleotide(job->forced[i])) : -1;
                T.nucFitchForwardPass(T.root, states, refState);
                T.nucFitchBackwardPass(T.root, states, (1 << getCodeFromNucleotide(job->consensus[i])));
                T.nucFitchAssignMutations(T.root, states, mutations, (1 << getCodeFromNucleotide(job->consensus[i])));
            } else {
                std::unordered_map<std::string, std::vector<int>> stateSets;
                std::unordered_map<std::string, int> states;
                for (const auto& u : *job->rows) {
                    std::vector<int> cur(16, SANKOFF_INF);
                    if (u.second[i] != '-') cur[getCodeFromNucleotide(u.second[i])] = 0;
                    else cur[0] = 0;
                    stateSets[u.first] = cur;
                }
                int defaultState = -1;
                if (job->forced) defaultState = (job->forced[i] != '-') ? getCodeFromNucleotide(job->forced[i]) : 0;
                T.nucSankoffForwardPass(T.root, stateSets);
                if (defaultState != -1)
                    T.nucSankoffBackwardPass(T.root, stateSets, states, getCodeFromNucleotide(job->consensus[i]), defaultState);
                else
                    T.nucSankoffBackwardPass(T.root, stateSets, states, getCodeFromNucleotide(job->consensus[i]));
                T.nucSankoffAssignMutations(T.root, states, mutations, getCodeFromNucleotide(job->consensus[i]));
            }
            for (const auto& m : mutations) {
                RefMut r{};
                r.node = job->graph->index.at(m.first);
                r.pos = (int32_t)i;
                r.type = (uint8_t)m.second.first;
                r.code = (uint8_t)getCodeFromNucleotide(m.second.second);
                local.push_back(r);
            }
        }
    }
    std::lock_guard<std::mutex> lk(job->outMutex);
    job->out->insert(job->out->end(), local.begin(), local.end());
}

void* threadMain(void* p) {
    runColumns(static_cast<ColumnJob*>(p));
    return nullptr;
}

size_t maxLevel(const NodeGraph& g) {
    size_t m = 1;
    for (auto* n : g.nodes) m = std::max(m, n->level);
    return m;
}

// Wall time of the last column loop (threads + the final sort), excluding the construction of the Node graph and of the
// row table around it -- those are one-off set-up in the reference too (createTreeFromNewickString, the FASTA read).
double g_lastLoopSeconds = 0;

// Runs the column loop on `nThreads` big-stack threads; output sorted by (node, pos) == per-node std::sort order.
void runAllColumns(ColumnJob& job, int nThreads, size_t depth) {
    const auto loopT0 = std::chrono::steady_clock::now();
    if (nThreads < 1) nThreads = 1;
    job.chunk = std::max<int64_t>(1, std::min<int64_t>(16, job.nCols / (4 * (int64_t)nThreads)));
    size_t stackBytes = (size_t)64 << 20;
    stackBytes += depth * 2048;  // Sankoff frames hold vector<vector<int>>; generous per-level allowance
    std::vector<pthread_t> th(nThreads);
    pthread_attr_t attr;
    pthread_attr_init(&attr);
    pthread_attr_setstacksize(&attr, stackBytes);
    for (int t = 0; t < nThreads; t++) pthread_create(&th[t], &attr, threadMain, &job);
    for (int t = 0; t < nThreads; t++) pthread_join(th[t], nullptr);
    pthread_attr_destroy(&attr);
    std::sort(job.out->begin(), job.out->end(), [](const RefMut& a, const RefMut& b) {
        return a.node != b.node ? a.node < b.node : a.pos < b.pos;
    });
    g_lastLoopSeconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - loopT0).count();
}

bool anyLeafPresent(const NodeGraph& g, const std::vector<std::pair<std::string, const char*>>& rows) {
    std::set<std::string> ids;
    for (auto& r : rows) ids.insert(r.first);
    for (auto* n : g.nodes)
        if (n->children.empty() && ids.count(n->identifier)) return true;
    return false;
}

// sort + <=6-run packing: src/panman.cpp:1435-1456 with the NucMut range ctor of src/panman.hpp:86-128.
// `muts` is sorted by (node,pos); emits per-node NucMut lists as one CSR.
void packNucMuts(const std::vector<RefMut>& muts, int nNodes, std::vector<int64_t>& off, std::vector<RefNucMut>& packed) {
    off.assign(nNodes + 1, 0);
    size_t a = 0;
    for (int node = 0; node < nNodes; node++) {
        off[node] = (int64_t)packed.size();
        size_t b = a;
        while (b < muts.size() && muts[b].node == node) b++;
        if (b > a) {
            std::vector<std::tuple<int, int8_t, int8_t>> v;
            for (size_t k = a; k < b; k++) v.emplace_back(muts[k].pos, (int8_t)muts[k].type, (int8_t)muts[k].code);
            std::sort(v.begin(), v.end());
            auto emit = [&](size_t start, size_t end) {
                RefNucMut m{};
                m.primaryBlockId = 0;
                m.secondaryBlockId = -1;
                m.mutInfo = (uint8_t)(((end - start) << 4) + (uint8_t)std::get<1>(v[start]));
                m.nucPosition = std::get<0>(v[start]);
                m.nucGapPosition = -1;
                m.nucs = 0;
                for (size_t i = start; i < end; i++) m.nucs += (std::get<2>(v[i]) << (4 * (5 - (i - start))));
                packed.push_back(m);
            };
            size_t currentStart = 0;
            for (size_t i = 1; i < v.size(); i++) {
                if (i - currentStart == 6 || std::get<0>(v[i]) != std::get<0>(v[i - 1]) + 1 ||
                    std::get<1>(v[i]) != std::get<1>(v[i - 1])) {
                    emit(currentStart, i);
                    currentStart = i;
                }
            }
            emit(currentStart, v.size());
        }
        a = b;
    }
    off[nNodes] = (int64_t)packed.size();
}

template <class T>
T* dupArray(const std::vector<T>& v) {
    T* p = (T*)malloc(std::max<size_t>(1, v.size()) * sizeof(T));
    if (!v.empty()) memcpy(p, v.data(), v.size() * sizeof(T));
    return p;
}

}  // namespace

extern "C" {

typedef RefMut pmref_mut;
typedef RefNucMut pmref_nucmut;

void pmref_free(void* p) { free(p); }
double pmref_last_loop_seconds(void) { return g_lastLoopSeconds; }

// Column loop over a flat tree.  Node i is named "node_<k>" if internal (k = 1 + number of internal nodes before it in
// pre-order, i.e. the reference's numbering), "r<row>" if leaf_row[i] >= 0, "absent_<i>" if leaf_row[i] == -2.  MSA rows
// are keyed "r<row>" and inserted in row order.  msa is [n_rows][n_cols] CHARACTERS; consensus/forced are characters.
// Returns 0, -4 (Sankoff with no tree leaf present in the MSA and no forced root: the reference asserts).
struct RefGraph {  // a flat tree turned into the reference's Node graph, kept across calls
    NodeGraph g;
    Tree tree;
    size_t depth = 1;
};

void* pmref_graph_new(int32_t n_nodes, const int32_t* parent, const int32_t* child_off, const int32_t* child_idx,
                      const int32_t* leaf_row) {
    RefGraph* R = new RefGraph();
    NodeGraph& g = R->g;
    g.nodes.resize(n_nodes);
    int internal = 0;
    for (int i = 0; i < n_nodes; i++) {
        Node* n = new Node();
        bool isLeaf = child_off[i + 1] == child_off[i];
        if (!isLeaf) n->identifier = "node_" + std::to_string(++internal);
        else if (leaf_row[i] >= 0) n->identifier = "r" + std::to_string(leaf_row[i]);
        else n->identifier = "absent_" + std::to_string(i);
        g.nodes[i] = n;
        g.index[n->identifier] = i;
    }
    for (int i = 0; i < n_nodes; i++) {
        Node* n = g.nodes[i];
        n->parent = parent[i] >= 0 ? g.nodes[parent[i]] : nullptr;
        for (int e = child_off[i]; e < child_off[i + 1]; e++) n->children.push_back(g.nodes[child_idx[e]]);
    }
    for (int i = 0; i < n_nodes; i++) {  // pre-order ids: parents precede children
        Node* n = g.nodes[i];
        n->level = n->parent ? n->parent->level + 1 : 1;
    }
    R->tree.root = g.nodes[0];
    R->depth = maxLevel(g);
    return R;
}

void pmref_graph_free(void* graph) { delete static_cast<RefGraph*>(graph); }

// The column loop on a graph built by pmref_graph_new (bench.py's reference arm calls it once per step).
int pmref_columns_g(void* graph, int algo, int32_t n_rows, int64_t n_cols, const char* msa, const char* consensus,
                    const char* forced, int n_threads, pmref_mut** out, int64_t* n_out) {
    RefGraph* R = static_cast<RefGraph*>(graph);
    std::vector<std::pair<std::string, const char*>> rows;
    for (int r = 0; r < n_rows; r++) rows.emplace_back("r" + std::to_string(r), msa + (size_t)r * n_cols);
    if (algo == 1 && !forced && !anyLeafPresent(R->g, rows)) return -4;
    std::vector<RefMut> muts;
    ColumnJob job;
    job.algo = algo;
    job.tree = &R->tree;
    job.graph = &R->g;
    job.rows = &rows;
    job.consensus = consensus;
    job.forced = forced;
    job.nCols = n_cols;
    job.out = &muts;
    runAllColumns(job, n_threads, R->depth);
    *out = dupArray(muts);
    *n_out = (int64_t)muts.size();
    return 0;
}

int pmref_columns(int algo, int32_t n_nodes, const int32_t* parent, const int32_t* child_off, const int32_t* child_idx,
                  const int32_t* leaf_row, int32_t n_rows, int64_t n_cols, const char* msa, const char* consensus,
                  const char* forced, int n_threads, pmref_mut** out, int64_t* n_out) {
    void* graph = pmref_graph_new(n_nodes, parent, child_off, child_idx, leaf_row);
    const int rc = pmref_columns_g(graph, algo, n_rows, n_cols, msa, consensus, forced, n_threads, out, n_out);
    pmref_graph_free(graph);
    return rc;
}

// Block-level passes (the per-block loop of the PanGraph branch, src/panman.cpp:865-950, restated; its callers build
// `alignedSequences` from a GFA / PanGraph, which is out of scope): one "column" per block, one state per sequence --
// 0 block absent, 1 forward strand, 2 reverse strand.  algo 0 = blockFitch*New (:898-904; a leaf without a row has the
// empty set), algo 1 = blockSankoff* (:943-949; a leaf without a row gets the blank state {0,INF,INF}).  forced: the
// reference sequence's states per block (defaultState) or nullptr.  Same tree/row naming as pmref_columns.
// Output records: node, pos = block, type = BlockMutationType (BI = 1, BD = 0), code = the inversion flag.
int pmref_block_columns(int algo, int32_t n_nodes, const int32_t* parent, const int32_t* child_off, const int32_t* child_idx,
                        const int32_t* leaf_row, int32_t n_rows, int64_t n_blocks, const uint8_t* states_in,
                        const uint8_t* forced, pmref_mut** out, int64_t* n_out) {
    NodeGraph g;
    g.nodes.resize(n_nodes);
    int internal = 0;
    for (int i = 0; i < n_nodes; i++) {
        Node* n = new Node();
        bool isLeaf = child_off[i + 1] == child_off[i];
        if (!isLeaf) n->identifier = "node_" + std::to_string(++internal);
        else if (leaf_row[i] >= 0) n->identifier = "r" + std::to_string(leaf_row[i]);
        else n->identifier = "absent_" + std::to_string(i);
        g.nodes[i] = n;
        g.index[n->identifier] = i;
    }
    for (int i = 0; i < n_nodes; i++) {
        Node* n = g.nodes[i];
        n->parent = parent[i] >= 0 ? g.nodes[parent[i]] : nullptr;
        for (int e = child_off[i]; e < child_off[i + 1]; e++) n->children.push_back(g.nodes[child_idx[e]]);
    }
    Tree tree;
    tree.root = g.nodes[0];
    std::vector<RefMut> muts;
    for (int64_t b = 0; b < n_blocks; b++) {
        std::unordered_map<std::string, std::pair<panmanUtils::BlockMutationType, bool>> mutations;
        if (algo == 0) {
            std::unordered_map<std::string, int> states;
            for (int r = 0; r < n_rows; r++) states["r" + std::to_string(r)] = 1 << states_in[(size_t)r * n_blocks + b];  // :888-896
            tree.blockFitchForwardPassNew(tree.root, states);
            if (forced) tree.blockFitchBackwardPassNew(tree.root, states, 1, 1 << forced[b]);  // :899-903
            else tree.blockFitchBackwardPassNew(tree.root, states, 1);
            tree.blockFitchAssignMutationsNew(tree.root, states, mutations, 1);
        } else {
            std::unordered_map<std::string, std::vector<int>> stateSets;
            std::unordered_map<std::string, int> states;
            for (int r = 0; r < n_rows; r++) {  // :931-941
                std::vector<int> cur(3, SANKOFF_INF);
                cur[states_in[(size_t)r * n_blocks + b]] = 0;
                stateSets["r" + std::to_string(r)] = cur;
            }
            tree.blockSankoffForwardPass(tree.root, stateSets);
            if (forced) tree.blockSankoffBackwardPass(tree.root, stateSets, states, 0, forced[b]);  // :944-948
            else tree.blockSankoffBackwardPass(tree.root, stateSets, states, 0);
            tree.blockSankoffAssignMutations(tree.root, states, mutations, 0);
        }
        for (const auto& m : mutations) {
            RefMut r = {};
            r.node = g.index.at(m.first);
            r.pos = (int32_t)b;
            r.type = (uint8_t)m.second.first;
            r.code = m.second.second ? 1 : 0;
            muts.push_back(r);
        }
    }
    std::sort(muts.begin(), muts.end(), [](const RefMut& a, const RefMut& b) { return a.node != b.node ? a.node < b.node : a.pos < b.pos; });
    *out = dupArray(muts);
    *n_out = (int64_t)muts.size();
    return 0;
}

// Whole construction path from files, as `panmanUtils -M msa -N nwk [--reference id] [--low-mem-mode]` runs it up to
// (not including) serialisation.  Outputs (malloc'd; release with pmref_free):
//   tree in pre-order (parent, child CSR, names as one '\n'-joined string), consensus characters,
//   (node,pos,type,code) records sorted by (node,pos), per-node packed NucMut CSR, Block consensus words.
// Returns 0, -1/-2 Newick errors, -3 unequal lengths, -4 Sankoff root invalid, -5 all-gap column without reference
// (undefined in the reference), -6 reference id absent (low-mem mode: the reference exits), -7 cannot open a file.
int pmref_build(const char* msa_path, const char* nwk_path, const char* reference_id, int low_mem, int n_threads,
                int32_t* n_nodes, int32_t** parent, int32_t** child_off, int32_t** child_idx, char** names_joined,
                int64_t* n_cols, char** consensus_out, pmref_mut** muts_out, int64_t* n_muts, int64_t** nuc_off,
                pmref_nucmut** nuc_out, int64_t* n_nuc, uint32_t** block_words, int64_t* n_block_words) {
    std::ifstream fin(msa_path), nwk(nwk_path);
    if (!fin.good() || !nwk.good()) return -7;
    std::string newickString;
    std::getline(nwk, newickString);
    NodeGraph g;
    int rc = newickToGraph(newickString, g);
    if (rc) return rc;
    Tree tree;
    tree.root = g.nodes[0];
    std::string reference = reference_id ? reference_id : "";

    std::map<std::string, std::string> seqs;
    size_t lineLength = 0;
    rc = low_mem ? readMsaLowMem(fin, seqs, lineLength) : readMsaDefault(fin, seqs, lineLength);
    if (rc) return rc;

    std::string consensusSeq;
    const char* forced = nullptr;
    if (!low_mem) {
        // src/panman.cpp:1322-1352
        if (reference != "") {
            consensusSeq = seqs[reference];  // operator[]: an absent id silently yields an empty consensus (zero columns)
        } else {
            consensusSeq.resize(lineLength);
            for (size_t i = 0; i < lineLength; i++) {
                bool found = false;
                for (auto& u : seqs)
                    if (u.second[i] != '-') {
                        consensusSeq[i] = u.second[i];
                        found = true;
                        break;
                    }
                if (!found) return -5;
            }
        }
        if (reference != "") forced = seqs[reference].c_str();
    } else {
        // src/panman.cpp:1517-1547
        consensusSeq.resize(lineLength);
        if (reference != "") {
            if (seqs.find(reference) == seqs.end()) return -6;
            const std::string& ref = seqs[reference];
            for (size_t i = 0; i < lineLength; i++) {
                if (ref[i] == '-') {
                    for (auto& u : seqs)
                        if (u.second[i] != '-') {
                            consensusSeq[i] = u.second[i];
                            break;
                        }
                } else {
                    consensusSeq[i] = ref[i];
                }
            }
            forced = ref.c_str();
        } else {
            for (size_t i = 0; i < lineLength; i++) {
                bool found = false;
                for (auto& u : seqs)
                    if (u.second[i] != '-') {
                        consensusSeq[i] = u.second[i];
                        found = true;
                        break;
                    }
                if (!found) return -5;  // reference: "ideally should not happen", exit(1)
            }
        }
    }

    std::vector<std::pair<std::string, const char*>> rows;
    for (auto& u : seqs) rows.emplace_back(u.first, u.second.c_str());
    if (low_mem && !forced && !anyLeafPresent(g, rows)) return -4;

    std::vector<RefMut> muts;
    ColumnJob job;
    job.algo = low_mem ? 1 : 0;
    job.tree = &tree;
    job.graph = &g;
    job.rows = &rows;
    job.consensus = consensusSeq.c_str();
    job.forced = forced;
    job.nCols = (int64_t)consensusSeq.size();
    job.out = &muts;
    runAllColumns(job, n_threads, maxLevel(g));

    int V = (int)g.nodes.size();
    std::vector<int32_t> par(V), coff(V + 1, 0), cidx;
    std::unordered_map<const Node*, int> idOf;
    for (int i = 0; i < V; i++) idOf[g.nodes[i]] = i;
    std::string names;
    for (int i = 0; i < V; i++) {
        par[i] = g.nodes[i]->parent ? idOf[g.nodes[i]->parent] : -1;
        coff[i] = (int32_t)cidx.size();
        for (auto* c : g.nodes[i]->children) cidx.push_back(idOf[c]);
        names += g.nodes[i]->identifier;
        names += '\n';
    }
    coff[V] = (int32_t)cidx.size();

    std::vector<int64_t> off;
    std::vector<RefNucMut> packed;
    packNucMuts(muts, V, off, packed);

    std::vector<uint32_t> words;  // Block(0, consensusSeq), src/panman.cpp:236-247
    for (size_t i = 0; i < consensusSeq.length(); i += 8) {
        uint32_t w = 0;
        for (size_t j = i; j < std::min(i + 8, consensusSeq.length()); j++) {
            int code = panmanUtils::getCodeFromNucleotide(consensusSeq[j]);
            w ^= (code << (4 * (7 - (j - i))));
        }
        words.push_back(w);
    }

    *n_nodes = V;
    *parent = dupArray(par);
    *child_off = dupArray(coff);
    *child_idx = dupArray(cidx);
    *names_joined = (char*)malloc(names.size() + 1);
    memcpy(*names_joined, names.c_str(), names.size() + 1);
    *n_cols = (int64_t)consensusSeq.size();
    *consensus_out = (char*)malloc(consensusSeq.size() + 1);
    memcpy(*consensus_out, consensusSeq.c_str(), consensusSeq.size() + 1);
    *muts_out = dupArray(muts);
    *n_muts = (int64_t)muts.size();
    *nuc_off = dupArray(off);
    *nuc_out = dupArray(packed);
    *n_nuc = (int64_t)packed.size();
    *block_words = dupArray(words);
    *n_block_words = (int64_t)words.size();
    return 0;
}

// Packing only (for checking a product packer against the reference rule on arbitrary records).
int pmref_pack(const pmref_mut* muts, int64_t n_muts, int32_t n_nodes, int64_t** nuc_off, pmref_nucmut** nuc_out,
               int64_t* n_nuc) {
    std::vector<RefMut> v(muts, muts + n_muts);
    std::vector<int64_t> off;
    std::vector<RefNucMut> packed;
    packNucMuts(v, n_nodes, off, packed);
    *nuc_off = dupArray(off);
    *nuc_out = dupArray(packed);
    *n_nuc = (int64_t)packed.size();
    return 0;
}

}  // extern "C"
