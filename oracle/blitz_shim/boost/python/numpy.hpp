// TEST INFRASTRUCTURE -- stand-in for <boost/python/numpy.hpp>: the reference's DGContext2D.hpp only DECLARES member
// functions returning these types (the pyblitzdg binding, out of scope); nothing of them is defined or called here.
#pragma once
namespace boost { namespace python {
class dict {};
namespace numpy { class ndarray {}; }
} }
