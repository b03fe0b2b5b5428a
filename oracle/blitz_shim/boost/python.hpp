// TEST INFRASTRUCTURE -- stand-in for <boost/python.hpp>: the reference's VtkOutputter.hpp only DECLARES member functions
// taking these types (the pyblitzdg binding, out of scope); nothing of them is defined or called here.
#pragma once
#include "python/numpy.hpp"
namespace boost { namespace python {
class str {};
} }
