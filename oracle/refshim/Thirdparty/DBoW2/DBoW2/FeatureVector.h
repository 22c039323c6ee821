/* oracle/refshim — TEST INFRASTRUCTURE.  Stand-in for Thirdparty/DBoW2/DBoW2/FeatureVector.h of the reference (which needs
 * boost::serialization, not in this image): the same container, `std::map<NodeId, std::vector<unsigned int>>`. */
#ifndef __D_T_FEATURE_VECTOR__
#define __D_T_FEATURE_VECTOR__
#include <map>
#include <vector>
namespace DBoW2 {
typedef unsigned int NodeId;
typedef unsigned int WordId;
class FeatureVector : public std::map<NodeId, std::vector<unsigned int> > {
public:
    void addFeature(NodeId id, unsigned int i_feature) { (*this)[id].push_back(i_feature); }
};
} // namespace DBoW2
#endif
