// minicv shim (TEST INFRASTRUCTURE): DBoW2's BowVector.h / FeatureVector.h befriend boost::serialization::access and
// carry an uninstantiated serialize() template (Thirdparty/DBoW2/DBoW2/BowVector.h:17-67); Boost is not installed and
// nothing on the BoW transform path serialises, so the two names are declared and nothing more.
#pragma once
namespace boost {
namespace serialization {
class access;
template <class Base, class Derived> Base& base_object(Derived& d) { return d; }
}  // namespace serialization
}  // namespace boost
