// transition_traits.h -- leaf order of the two record types of the reference
// (rela/types.h:18-73). Written against member NAMES only, so the same traits
// serve rela_b200's own structs (types.h) and the reference's rela::FFTransition
// / rela::RNNTransition when the drop-in header is used.
//
// Leaf order (fixed; it is the schema order of the slabs):
//   FF : obs.* (sorted), action.* (sorted), reward, terminal, bootstrap, next_obs.* (sorted)
//   RNN: obs.*, h0.*, action.*, reward, terminal, bootstrap, seq_len
#pragma once

#include "rela_b200/prioritized_replay.h"

namespace rela_b200 {

template <class T>
struct FFTraits {
  static constexpr bool kTimeMajor = false;

  static Leaves leaves(const T& t) {
    Leaves out;
    appendDict(out, "obs", t.obs);
    appendDict(out, "action", t.action);
    out.push_back(Leaf{"reward", t.reward});
    out.push_back(Leaf{"terminal", t.terminal});
    out.push_back(Leaf{"bootstrap", t.bootstrap});
    appendDict(out, "next_obs", t.nextObs);
    return out;
  }

  static T fromLeaves(Leaves& leaves, int /*batchsize*/) {
    T t;
    std::string key;
    for (Leaf& lf : leaves) {
      if (leafInGroup(lf.name, "obs", &key)) t.obs.insert({key, lf.t});
      else if (leafInGroup(lf.name, "action", &key)) t.action.insert({key, lf.t});
      else if (leafInGroup(lf.name, "next_obs", &key)) t.nextObs.insert({key, lf.t});
      else if (lf.name == "reward") t.reward = lf.t;
      else if (lf.name == "terminal") t.terminal = lf.t;
      else if (lf.name == "bootstrap") t.bootstrap = lf.t;
    }
    return t;
  }
};

template <class T>
struct RNNTraits {
  static constexpr bool kTimeMajor = true;

  static Leaves leaves(const T& t) {
    Leaves out;
    appendDict(out, "obs", t.obs);
    appendDict(out, "h0", t.h0);
    appendDict(out, "action", t.action);
    out.push_back(Leaf{"reward", t.reward});
    out.push_back(Leaf{"terminal", t.terminal});
    out.push_back(Leaf{"bootstrap", t.bootstrap});
    out.push_back(Leaf{"seq_len", t.seqLen});
    return out;
  }

  static T fromLeaves(Leaves& leaves, int batchsize) {
    T t;
    std::string key;
    for (Leaf& lf : leaves) {
      if (leafInGroup(lf.name, "obs", &key)) t.obs.insert({key, lf.t});
      else if (leafInGroup(lf.name, "h0", &key)) t.h0.insert({key, lf.t});
      else if (leafInGroup(lf.name, "action", &key)) t.action.insert({key, lf.t});
      else if (lf.name == "reward") t.reward = lf.t;
      else if (lf.name == "terminal") t.terminal = lf.t;
      else if (lf.name == "bootstrap") t.bootstrap = lf.t;
      else if (lf.name == "seq_len") t.seqLen = lf.t.reshape({batchsize});  // stack(0).squeeze(1), types.cc:167
    }
    return t;
  }
};

}  // namespace rela_b200
