/* KeyValue as seen by map()/reduce() callbacks: source-compatible with the public part of the reference's
 * class (src/keyvalue.h:23-59) -- the statistics fields and the three add() forms.  Written from scratch:
 * pairs added by a host callback are batched on the host and shipped to the HBM-resident store through
 * mrgpu_kv_add() (include/mrgpu.h). */
#ifndef KEY_VALUE_H
#define KEY_VALUE_H

#include "mpi.h"     // the reference's keyvalue.h brings these three in (src/keyvalue.h:17-19); callers rely on it
#include <stdio.h>
#include <stdint.h>
#include <vector>

struct mrgpu_mr;

namespace MAPREDUCE_NS {

class KeyValue {
  friend class MapReduce;

 public:
  uint64_t nkv;    // # of KV pairs in entire KV on this proc
  uint64_t ksize;  // exact size of all key data
  uint64_t vsize;  // exact size of all value data
  uint64_t esize;  // exact size of all data in KV
  uint64_t fsize;  // size of KV file (always 0: pages live in HBM)
  int msize;       // size of largest KV pair across all procs (not tracked: 0)
  char *page;      // in-memory page (host copy while a scan/map walks it, else NULL)
  int memtag;
  int npage;       // # of pages in entire KV

  void add(char *key, int keybytes, char *value, int valuebytes);               // src/keyvalue.cpp:343
  void add(int n, char *key, int keybytes, char *value, int valuebytes);        // src/keyvalue.cpp:399
  void add(int n, char *key, int *keybytes, char *value, int *valuebytes);      // src/keyvalue.cpp:419

  // Callers that feed one MapReduce's KeyValue from a callback running on ANOTHER one bracket the adds themselves
  // (oink/sssp.cpp:112-137: mrvert->kv->append(); mrpath->map(mrpath, move_to_new_mr, mrvert); mrvert->kv->complete();)
  void append();    // src/keyvalue.cpp:185-209: re-open for more pairs
  void complete();  // src/keyvalue.cpp:216-255: finish; the statistics fields above are current afterwards

 private:
  explicit KeyValue(mrgpu_mr *h);
  void flush();
  mrgpu_mr *handle;  // the handle pairs are shipped to
  std::vector<char> kbuf, vbuf;
  std::vector<int32_t> klen, vlen;
};

}  // namespace MAPREDUCE_NS

#endif
