// TEST INFRASTRUCTURE (oracle/_ref build only): Lua is not in this image; the dictionary header
// only needs the state type's name.
#ifndef ORACLE_REF_SHIM_LUA_HPP_
#define ORACLE_REF_SHIM_LUA_HPP_
struct lua_State;
#endif
