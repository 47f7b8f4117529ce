// TEST INFRASTRUCTURE ONLY (oracle/_ref build).
//
// Driver around the UNMODIFIED reference trie + clustering code: it #includes
// lib/trie/src/trie.tcc (via src/leaf.h) and src/cluster.cc from /root/reference where
// they lie, repeats the three loops of src/humid.cc (:117-126 / :145-154 neighbours,
// :178-189 clusters) verbatim in spirit, and prints what the `humid` binary never prints:
// every leaf's neighbour list (order and multiplicity) and cluster assignment.
//
// stdin:   line 1: "<H|E> <distance> <X|D>"   (Hamming|Edit, -m, maximum|directional)
//          then one word per line: "<ACGT...> [count]"   (count defaults to 1)
// stdout:  "W <rank> <word> <count> <cluster id> <n> <nbr rank>*n"   per leaf, walk order
//          "C <id> <size> <maxLeaf rank>"                            per cluster
//          "Q <word> <n> <result word>*n"  for optional query lines "?<h|H|l|L> <d> <word>"
//             (h = hamming, H = asymmetricHamming, l = levenshtein, L = asymmetricLevenshtein)
#include <cstdio>
#include <iostream>
#include <map>
#include <sstream>
#include <string>

#include "src/cluster.cc"  // pulls cluster.h, leaf.h -> lib/trie/src/trie.tcc

using namespace std;

static vector<uint8_t> encode(string const& s) {
  vector<uint8_t> w;
  for (char c : s) w.push_back(c == 'A' ? 0 : c == 'C' ? 1 : c == 'G' ? 2 : 3);
  return w;
}
static string decode(vector<uint8_t> const& w) {
  string s;
  for (uint8_t c : w) s.push_back("ACGT"[c]);
  return s;
}

int main() {
  string mode, clus;
  int dist;
  string line;
  getline(cin, line);
  {
    istringstream is(line);
    is >> mode >> dist >> clus;
  }
  Trie<4, NLeaf> trie;
  vector<string> queries;
  while (getline(cin, line)) {
    if (line.empty()) continue;
    if (line[0] == '?') { queries.push_back(line); continue; }
    istringstream is(line);
    string w; size_t c = 1;
    is >> w; is >> c;
    for (size_t i = 0; i < c; i++) trie.add(encode(w));
  }
  for (string const& q : queries) {
    istringstream is(q.substr(1));
    string kind, w; int d;
    is >> kind >> d >> w;
    vector<string> out;
    vector<uint8_t> word = encode(w);
    if (kind == "h") for (auto const& r : trie.hamming(word, d)) out.push_back(decode(r.path));
    if (kind == "H") for (auto const& r : trie.asymmetricHamming(word, d)) out.push_back(decode(r.path));
    if (kind == "l") for (auto const& r : trie.levenshtein(word, d)) out.push_back(decode(r.path));
    if (kind == "L") for (auto const& r : trie.asymmetricLevenshtein(word, d)) out.push_back(decode(r.path));
    printf("Q %s %zu", w.c_str(), out.size());
    for (auto const& s : out) printf(" %s", s.empty() ? "-" : s.c_str());
    printf("\n");
  }

  map<NLeaf*, size_t> rank;
  size_t n = 0;
  for (Result<NLeaf> const& r : trie.walk()) rank[r.leaf] = n++;

  // src/humid.cc:117-126 (Hamming) and :145-154 (edit).
  for (Result<NLeaf> const& w : trie.walk()) {
    if (mode == "H") {
      for (Result<NLeaf> const& r : trie.asymmetricHamming(w.path, dist))
        if (w.leaf != r.leaf) { w.leaf->neighbours.push_back(r.leaf); r.leaf->neighbours.push_back(w.leaf); }
    } else {
      for (Result<NLeaf> const& r : trie.asymmetricLevenshtein(w.path, dist))
        if (w.leaf != r.leaf) { w.leaf->neighbours.push_back(r.leaf); r.leaf->neighbours.push_back(w.leaf); }
    }
  }
  // src/humid.cc:178-189.
  vector<Cluster*> clusters;
  size_t id = 1;
  for (Result<NLeaf> const& r : trie.walk()) {
    if (!r.leaf->cluster) {
      Cluster* c = new Cluster{id++};
      if (clus == "X") assignMaxCluster(r.leaf, c); else assignDirectionalCluster(r.leaf, c);
      clusters.push_back(c);
    }
  }
  for (Result<NLeaf> const& r : trie.walk()) {
    printf("W %zu %s %zu %zu %zu", rank[r.leaf], decode(r.path).c_str(), r.leaf->count,
           r.leaf->cluster->id, r.leaf->neighbours.size());
    for (NLeaf* nb : r.leaf->neighbours) printf(" %zu", rank[nb]);
    printf("\n");
  }
  for (Cluster* c : clusters) printf("C %zu %zu %zu\n", c->id, c->size, rank[c->maxLeaf]);
  freeClusters(clusters);
  return 0;
}
