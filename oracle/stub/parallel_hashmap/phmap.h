// TEST INFRASTRUCTURE ONLY (oracle).  The reference includes
// "parallel_hashmap/phmap.h" at csrc/cpu/utils.h:4 but uses no symbol from it
// (SURVEY.md section 2.2), so an empty header is a faithful stand-in.
#pragma once
