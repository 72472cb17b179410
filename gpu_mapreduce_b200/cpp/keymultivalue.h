/* KeyMultiValue as callers see it.  The reference's class (src/keymultivalue.h) is internal to the library: callers
 * include the header (oink/sssp.cpp does) and test MapReduce::kmv for NULL, nothing more.  Here the KMV lives in HBM
 * behind the C-ABI handle; this type only gives that pointer something to point at. */
#ifndef KEY_MULTIVALUE_H
#define KEY_MULTIVALUE_H

namespace MAPREDUCE_NS {
class KeyMultiValue {};
}  // namespace MAPREDUCE_NS

#endif
