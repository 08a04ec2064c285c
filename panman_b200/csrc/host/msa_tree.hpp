// Host C++ mirror of the reference's MSA construction path, with the per-column loop replaced by one libpmfs call.
//
// Mirrors (paths relative to /root/reference):
//   panmanUtils::Tree::Tree(std::ifstream& fin, std::ifstream& secondFin, FILE_TYPE ftype, std::string reference)
//       src/panman.hpp:458-459, branches FILE_TYPE::MSA (src/panman.cpp:1264-1456) and MSA_OPTIMIZE (:1457-1639)
//   createTreeFromNewickString  src/panman.cpp:300-440        stringSplit / stripString  src/panman.cpp:255-298
//   code tables                 src/panman.cpp:46-118         NucMut range ctor          src/panman.hpp:86-128
//   Block(0, consensusSeq)      src/panman.cpp:236-247
// Same names, argument meaning and results; where the reference exit()s or asserts, BuildError is thrown instead.
#pragma once

#include <cstdint>
#include <istream>
#include <map>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../../include/pmfs.h"

namespace pmfs_host {

enum FILE_TYPE { PANMAT = 0, GFA = 1, PANGRAPH = 2, MSA = 3, MSA_OPTIMIZE = 4 };  // src/common.hpp:36-43
enum NucMutationType { NS = 0, ND = 1, NI = 2 };                                 // src/panman.hpp:26-41 (MSA path subset)

struct BuildError : std::runtime_error {
    int status;  // pmfs_status, or -10.. for host-side input errors
    BuildError(int s, const std::string& m) : std::runtime_error(m), status(s) {}
};

char getNucleotideFromCode(int code);
char getCodeFromNucleotide(char nuc);
void stringSplit(std::string const& s, char delim, std::vector<std::string>& words);
std::string stripString(std::string s);

struct NucMut {  // src/panman.hpp:55-61
    int32_t nucPosition;
    int32_t nucGapPosition;
    int32_t primaryBlockId;
    int32_t secondaryBlockId;
    uint8_t mutInfo;
    uint32_t nucs;
    NucMut() = default;
    // "Create non-SNP mutations from SNP mutations at consecutive positions for MSA" (src/panman.hpp:86-128) on records
    NucMut(const pmfs_mut* mutationArray, int start, int end);
};

struct Node {  // src/panman.hpp:307-322, children/parent as indices into Tree::nodes
    float branchLength = 0.0;
    size_t level = 1;
    std::string identifier;
    int32_t parent = -1;
    std::vector<int32_t> children;
    std::vector<NucMut> nucMutation;
    bool blockInsertionAtRoot = false;  // root->blockMutation.emplace_back(0, {BI,false}) (src/panman.cpp:1430)
};

class Tree {
  public:
    std::vector<Node> nodes;                         // creation order == pre-order; nodes[0] is the root
    std::unordered_map<std::string, int32_t> allNodes;
    std::string consensusSeq;
    std::vector<uint32_t> blockConsensusSeq;          // blocks[0].consensusSeq
    std::vector<pmfs_mut> mutations;                  // all records, sorted by (node,pos)
    pmfs_timing timing = {};

    // Same signature as the reference constructor plus the CUDA device to run on.
    Tree(std::istream& fin, std::istream& secondFin, FILE_TYPE ftype, std::string reference = "", int device = 0);
    // An empty tree for code that fills the fields itself (the writer's array entry point).
    explicit Tree(std::nullptr_t) {}

    // Pieces, exposed so each can be tested against the reference behaviour on its own
    static void createTreeFromNewickString(std::string newick, std::vector<Node>& nodes,
                                           std::unordered_map<std::string, int32_t>& allNodes);
    static void readMSA(std::istream& fin, FILE_TYPE ftype, std::map<std::string, std::string>& sequenceIdsToSequences,
                        size_t& lineLength);
    static std::string computeConsensus(const std::map<std::string, std::string>& seqs, size_t lineLength, FILE_TYPE ftype,
                                        const std::string& reference);
    static void packNucMutations(const std::vector<pmfs_mut>& sortedMuts, std::vector<Node>& nodes);
    static std::vector<uint32_t> packBlockConsensus(const std::string& seq);
};

// ---- .panman output (panman_writer.cpp) ----
// Tree::getNewickString (src/panman.cpp:1923-2030) of the whole tree.
std::string getNewickString(const Tree& tree);
// The Cap'n Proto stream TreeGroup::writeToFile (src/panman.cpp:7027-7161) produces for a TreeGroup holding this tree.
std::vector<uint8_t> encodePanMAN(const Tree& tree);
// writePanMAN (src/panmanUtils.cpp:269-297): the stream through LZMA level 9 (xz) into `path`; compress = false leaves
// it bare.
void writePanMAN(const Tree& tree, const std::string& path, bool compress = true);

}  // namespace pmfs_host
