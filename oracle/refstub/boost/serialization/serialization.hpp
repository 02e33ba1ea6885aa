/* Stub: DBoW2 (Thirdparty/DBoW2/DBoW2/BowVector.h:17-18, FeatureVector.h:18-19) only names boost::serialization in member templates that
 * are never instantiated on the path under test.  Test infrastructure (oracle/_ref build). */
#ifndef ORBFE_BOOST_SER_STUB
#define ORBFE_BOOST_SER_STUB
namespace boost { namespace serialization {
class access;
template <class Base, class Derived> Base& base_object(Derived& d) { return static_cast<Base&>(d); }
} }
#endif
