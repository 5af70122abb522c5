// KD_Tree parameters (engine/renderer/kd_tree.h:13-47). The node array lives inside the model handle
// (atray_model); `tree.length` mirrors the node count so prep_scene's "build only if empty" test keeps working.
#pragma once
#include "PL/PL_math.h"
enum class KD_Division_Method { CENTER, SAH };
struct KD_NodeArrayView { int32 length = 0; const void* front = nullptr; };
struct KD_Tree
{
    uint32 max_no_faces_per_node = 300;
    KD_Division_Method division_method = KD_Division_Method::SAH;
    KD_NodeArrayView tree;
};
