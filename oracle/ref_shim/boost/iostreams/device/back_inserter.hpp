// TEST INFRASTRUCTURE (oracle/_ref build only): see boost/iostreams/filtering_stream.hpp.
#include "boost/iostreams/filtering_stream.hpp"
