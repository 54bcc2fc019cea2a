// oracle/ref_shim -- TEST INFRASTRUCTURE.  The reference names
// boost::multiprecision::cpp_int only inside a static_assert
// (include/utils/next_combmask.h:38-40); a declaration is all it needs.
#pragma once
namespace boost {
namespace multiprecision {
struct cpp_int;
}
}  // namespace boost
