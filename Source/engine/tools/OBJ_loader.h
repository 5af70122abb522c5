// load_model_data (engine/tools/OBJ_loader.h:6, OBJ_loader.cpp:278-361): threadpool.threads.size = number of parse chunks.
#pragma once
#include "engine/renderer/model.h"
#include "thread_pool.h"
inline void load_model_data(ModelData& mdl, const char* file_name, ThreadPool& threadpool)
{
    atray_model_load_obj(file_name, threadpool.threads.size > 0 ? threadpool.threads.size : 1, &mdl.handle);
}
