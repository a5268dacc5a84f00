// loader.h -- scene XML loading (reference: SceneHandler, include/mitsuba/render/scenehandler.h;
// Utility::loadScene, src/librender/util.cpp).
#pragma once
#include <map>
#include <memory>
#include <string>

#include "plugins.h"

namespace mtsbhost {

typedef std::map<std::string, std::string> ParameterMap;      // -Dname=value command line definitions

// Parses `xml` (fileName is used for messages and to resolve relative mesh / include paths), builds
// and configures the object graph and returns the Scene.  Errors are thrown as std::runtime_error
// with the reference's "Error in file ... (near line N): ..." form.
std::shared_ptr<Scene> loadSceneString(const std::string &xml, const std::string &fileName, ParameterMap &params,
                                       std::map<std::string, ObjRef> *namedObjects = nullptr, bool included = false);
std::shared_ptr<Scene> loadSceneFile(const std::string &path, ParameterMap &params,
                                     std::map<std::string, ObjRef> *namedObjects = nullptr, bool included = false);

}  // namespace mtsbhost
