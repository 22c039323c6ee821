/* oracle/boostshim — TEST INFRASTRUCTURE.  The names the reference's vendored DBoW2 headers mention from Boost.Serialization
 * (Thirdparty/DBoW2/DBoW2/BowVector.h:17-18,62-66, FeatureVector.h:18-19,27-31).  The serialize() member templates that use
 * them are never instantiated by the oracle (nothing is saved or loaded through Boost), so declarations are enough. */
#pragma once
namespace boost { namespace serialization {
class access;
template <class Base, class Derived> Base& base_object(Derived& d) { return static_cast<Base&>(d); }
template <class T> T& make_array(T*, unsigned long);
} }
