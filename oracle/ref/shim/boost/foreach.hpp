// stand-in for <boost/foreach.hpp> (Boost is absent); reference uses BOOST_FOREACH as a range-for
#pragma once
#define BOOST_FOREACH(decl, cont) for (decl : cont)
