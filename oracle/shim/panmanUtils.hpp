// TEST INFRASTRUCTURE ONLY (oracle). Not part of the product; nothing under panman_b200/ may include this.
//
// Minimal stand-in for the reference's "panmanUtils.hpp" so that the reference's own
// src/fitchSankoff.cpp can be compiled UNMODIFIED, in place, from /root/reference (see oracle/Makefile:
// the file is fed to g++ on stdin with this directory as the working directory, so its
// `#include "panmanUtils.hpp"` resolves here instead of to the real header, which needs TBB, Boost,
// Cap'n Proto, protobuf and jsoncpp -- none of which exist in this image).
//
// Only declarations are provided; every definition of the 18 Fitch/Sankoff members comes from the
// reference translation unit.  Declarations follow /root/reference/src/panman.hpp:26-52 (enums),
// :307-322 (Node), :473-529 (member prototypes incl. default arguments) and src/common.hpp:16 (SANKOFF_INF).
#pragma once

#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstdint>
#include <iostream>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

static const int SANKOFF_INF = 100000001;  // src/common.hpp:16

namespace panmanUtils {

enum NucMutationType { NS = 0, ND = 1, NI = 2, NSNPS = 3, NSNPI = 4, NSNPD = 5, NNONE = 2000 };  // panman.hpp:26-41
enum BlockMutationType { BI = 1, BD = 0, BIn = 2, NONE = 1000 };                                // panman.hpp:43-52

char getNucleotideFromCode(int code);   // defined in oracle/ref_driver.cpp, restating src/panman.cpp:46-81
char getCodeFromNucleotide(char nuc);   // restating src/panman.cpp:83-118

class Node {  // the fields fitchSankoff.cpp touches (panman.hpp:307-322)
  public:
    float branchLength = 0.0;
    size_t level = 1;
    std::string identifier;
    Node* parent = nullptr;
    std::vector<Node*> children;
};

class Tree {
  public:
    Node* root = nullptr;

    int nucFitchForwardPass(Node* node, std::unordered_map<std::string, int>& states, int refState = -1);
    int nucFitchForwardPassOpt(Node* node, std::unordered_map<std::string, int>& states);
    void nucFitchBackwardPass(Node* node, std::unordered_map<std::string, int>& states, int parentState,
                              int defaultState = (1 << 28));
    void nucFitchBackwardPassOpt(Node* node, std::unordered_map<std::string, int>& states, int parentState,
                                 int defaultState = (1 << 28));
    void nucFitchAssignMutations(Node* node, std::unordered_map<std::string, int>& states,
                                 std::unordered_map<std::string, std::pair<panmanUtils::NucMutationType, char>>& mutations,
                                 int parentState);
    void nucFitchAssignMutationsOpt(Node* node, std::unordered_map<std::string, int>& states,
                                    std::unordered_map<std::string, std::pair<panmanUtils::NucMutationType, char>>& mutations,
                                    int parentState);

    std::vector<int> nucSankoffForwardPass(Node* node, std::unordered_map<std::string, std::vector<int>>& stateSets);
    std::vector<int> nucSankoffForwardPassOpt(Node* node, std::unordered_map<std::string, std::vector<int>>& stateSets);
    void nucSankoffBackwardPass(Node* node, std::unordered_map<std::string, std::vector<int>>& stateSets,
                                std::unordered_map<std::string, int>& states, int parentPtr, int defaultValue = (1 << 28));
    void nucSankoffBackwardPassOpt(Node* node, std::unordered_map<std::string, std::vector<int>>& stateSets,
                                   std::unordered_map<std::string, int>& states, int parentPtr,
                                   int defaultValue = (1 << 28));
    void nucSankoffAssignMutations(Node* node, std::unordered_map<std::string, int>& states,
                                   std::unordered_map<std::string, std::pair<panmanUtils::NucMutationType, char>>& mutations,
                                   int parentState);
    void nucSankoffAssignMutationsOpt(Node* node, std::unordered_map<std::string, int>& states,
                                      std::unordered_map<std::string, std::pair<panmanUtils::NucMutationType, char>>& mutations,
                                      int parentState);

    int blockFitchForwardPassNew(Node* node, std::unordered_map<std::string, int>& states);
    void blockFitchBackwardPassNew(Node* node, std::unordered_map<std::string, int>& states, int parentState,
                                   int defaultValue = (1 << 28));
    void blockFitchAssignMutationsNew(Node* node, std::unordered_map<std::string, int>& states,
                                      std::unordered_map<std::string, std::pair<panmanUtils::BlockMutationType, bool>>& mutations,
                                      int parentState);

    std::vector<int> blockSankoffForwardPass(Node* node, std::unordered_map<std::string, std::vector<int>>& stateSets);
    void blockSankoffBackwardPass(Node* node, std::unordered_map<std::string, std::vector<int>>& stateSets,
                                  std::unordered_map<std::string, int>& states, int parentPtr, int defaultValue = (1 << 28));
    void blockSankoffAssignMutations(Node* node, std::unordered_map<std::string, int>& states,
                                     std::unordered_map<std::string, std::pair<panmanUtils::BlockMutationType, bool>>& mutations,
                                     int parentState);
};

}  // namespace panmanUtils
