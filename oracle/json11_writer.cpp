// TEST INFRASTRUCTURE -- the reference's JSON writer for the scene / motion files (OptiMo/core_io.cpp:36-98,216-226,252-294), over
// the reference's OWN json11 (external/json11/json11.cpp, compiled where it lies by `make -C oracle ref`). OptiMo's core_io.cpp
// itself needs Qt + Eigen and cannot be compiled here; what it does with json11 is restated below, call for call, so that the byte
// format (key order = std::map order, numbers through "%.17g" / "%d", ", " / ": " separators) is json11's, not a guess.
//
//   json11_writer scene <scene.json>     re-emits a scene file as Core::ConvertCurrentSceneToJson would after reading it
//   json11_writer canon                  reads JSON text on stdin, writes Json::parse(text).dump(): the canonical json11 form
#include <fstream>
#include <iostream>
#include <iterator>
#include <sstream>

#include "json11.hpp"

using json11::Json;

static Json vec(const Json& a)   // InterpretVector3d2Json / InterpretVectorXd2Json: every entry a double
{
    std::vector<Json> v;
    for (const Json& e : a.array_items()) v.push_back(Json(e.number_value()));
    return v;
}
static Json variable(const Json& j)   // InterpretVariable2Json, core_io.cpp:49-56
{
    std::vector<Json> pts;
    for (const Json& p : j["curve"].array_items()) pts.push_back(vec(p));
    return Json::object{{"var", Json(j["var"].number_value())}, {"curve", pts}};
}
static Json joint(const Json& j)      // InterpretJoint2Json, core_io.cpp:58-75
{
    std::vector<Json> children;
    for (const Json& c : j["children"].array_items()) children.push_back(joint(c));
    return Json::object{{"name", Json(j["name"].string_value())}, {"offset", vec(j["offset"])}, {"rot_x", variable(j["rot_x"])},
                        {"rot_y", variable(j["rot_y"])}, {"rot_z", variable(j["rot_z"])}, {"children", children}};
}
static Json ik(const Json& j)         // InterpretIkHandler2Json, core_io.cpp:77-88
{
    return Json::object{{"name", Json(j["name"].string_value())}, {"pos_x", variable(j["pos_x"])}, {"pos_y", variable(j["pos_y"])},
                        {"pos_z", variable(j["pos_z"])}, {"root_name", Json(j["root_name"].string_value())},
                        {"end_effector_name", Json(j["end_effector_name"].string_value())}};
}

int main(int argc, char** argv)
{
    const std::string mode = argc > 1 ? argv[1] : "";
    std::string text, err;
    if (mode == "scene" && argc > 2) {
        std::ifstream in(argv[2]);
        text.assign(std::istreambuf_iterator<char>(in), std::istreambuf_iterator<char>());
        const Json s = Json::parse(text, err);
        if (!err.empty()) { std::cerr << err << "\n"; return 1; }
        std::vector<Json> iks;
        for (const Json& h : s["ik_handlers"].array_items()) iks.push_back(ik(h));
        const Json obj = Json::object{{"min_frame", Json(s["min_frame"].int_value())}, {"max_frame", Json(s["max_frame"].int_value())},   // core_io.cpp:216-226
                                      {"frames_per_second", Json(s["frames_per_second"].int_value())}, {"joints", joint(s["joints"])},
                                      {"ik_handlers", iks}};
        std::cout << obj.dump();
        return 0;
    }
    if (mode == "canon") {
        text.assign(std::istreambuf_iterator<char>(std::cin), std::istreambuf_iterator<char>());
        const Json s = Json::parse(text, err);
        if (!err.empty()) { std::cerr << err << "\n"; return 1; }
        std::cout << s.dump();
        return 0;
    }
    std::cerr << "usage: json11_writer scene <file> | canon < text\n";
    return 2;
}
