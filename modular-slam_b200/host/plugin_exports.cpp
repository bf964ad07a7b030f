// plugin_exports.cpp -- factory functions + boost::dll aliases + the plugin_config.json reader.
//
// The reference loads plugins with
//   boost::dll::import_alias<std::unique_ptr<BlockType>()>(lib, name, load_mode::append_decorations)   plugin_loader.hpp:19-25
// and a plugin publishes a factory with BOOST_DLL_ALIAS(function, aliasName)                             test/dummy.cpp:11
// which expands (boost/dll/alias.hpp) to an exported, weak `extern "C" const void* aliasName` placed in section
// "boostdll" that holds the address of the function.  Boost headers are not available in this image, so the alias is
// spelled out; with Boost present, -DMSLAM_B200_WITH_BOOST uses the real macro.
#include <cctype>
#include <cstdlib>
#include <map>
#include <memory>

#include "orb_b200_plugin.hpp"
#include "ransac_pnp_b200.hpp"

#ifdef MSLAM_B200_WITH_BOOST
#include <boost/dll/alias.hpp>
#define MSLAM_B200_DLL_ALIAS(fn, alias) BOOST_DLL_ALIAS(fn, alias)
#else
#define MSLAM_B200_DLL_ALIAS(fn, alias)                                                                                \
    namespace _autoaliases                                                                                             \
    {                                                                                                                  \
    extern "C" __attribute__((visibility("default"))) const void* alias;                                              \
    __attribute__((section("boostdll"), weak)) const void* alias =                                                     \
        reinterpret_cast<const void*>(reinterpret_cast<std::intptr_t>(&fn));                                           \
    }
#endif

namespace mslam
{

std::unique_ptr<IOrbFeatureDetector> createOrbB200Detector()
{
    return std::unique_ptr<IOrbFeatureDetector>(new OrbB200Detector());
}

std::unique_ptr<IOrbFeatureDetector> createDistributedOrbB200Detector()
{
    return std::unique_ptr<IOrbFeatureDetector>(new DistributedOrbB200Detector());
}

std::unique_ptr<IRansacPnp> createRansacPnpB200(); // ransac_pnp_b200.cpp

std::unique_ptr<IOrbMatcher> createOrbB200Matcher()
{
    return std::unique_ptr<IOrbMatcher>(new OrbB200Matcher());
}

// ---- a small JSON reader (objects, arrays, strings, numbers, true/false/null) --------------------------------------
namespace
{
struct JsonValue
{
    enum Kind { Null, Bool, Number, String, Array, Object } kind = Null;
    double number = 0;
    bool boolean = false;
    std::string string;
    std::vector<JsonValue> array;
    std::vector<std::pair<std::string, JsonValue>> object;
    const JsonValue* find(const std::string& key) const
    {
        for(const auto& kv : object)
            if(kv.first == key)
                return &kv.second;
        return nullptr;
    }
};

struct JsonParser
{
    const std::string& s;
    size_t i = 0;
    std::string error;
    explicit JsonParser(const std::string& text) : s(text) {}
    void ws()
    {
        while(i < s.size() && std::isspace((unsigned char)s[i]))
            ++i;
    }
    bool fail(const char* what)
    {
        if(error.empty())
            error = std::string(what) + " at offset " + std::to_string(i);
        return false;
    }
    bool parseString(std::string* out)
    {
        if(i >= s.size() || s[i] != '"')
            return fail("expected string");
        ++i;
        out->clear();
        while(i < s.size() && s[i] != '"')
        {
            if(s[i] == '\\')
            {
                if(++i >= s.size())
                    return fail("bad escape");
                switch(s[i])
                {
                case 'n': out->push_back('\n'); break;
                case 't': out->push_back('\t'); break;
                case 'r': out->push_back('\r'); break;
                case 'b': out->push_back('\b'); break;
                case 'f': out->push_back('\f'); break;
                case 'u':
                    if(i + 4 >= s.size())
                        return fail("bad \\u escape");
                    out->push_back((char)std::strtol(s.substr(i + 1, 4).c_str(), nullptr, 16)); // ASCII subset
                    i += 4;
                    break;
                default: out->push_back(s[i]);
                }
                ++i;
            }
            else
                out->push_back(s[i++]);
        }
        if(i >= s.size())
            return fail("unterminated string");
        ++i;
        return true;
    }
    bool parseValue(JsonValue* v, int depth = 0)
    {
        if(depth > 32)
            return fail("nesting too deep");
        ws();
        if(i >= s.size())
            return fail("unexpected end");
        const char c = s[i];
        if(c == '{')
        {
            v->kind = JsonValue::Object;
            ++i;
            ws();
            if(i < s.size() && s[i] == '}')
                return ++i, true;
            for(;;)
            {
                ws();
                std::string key;
                if(!parseString(&key))
                    return false;
                ws();
                if(i >= s.size() || s[i] != ':')
                    return fail("expected ':'");
                ++i;
                JsonValue child;
                if(!parseValue(&child, depth + 1))
                    return false;
                v->object.emplace_back(std::move(key), std::move(child));
                ws();
                if(i < s.size() && s[i] == ',')
                {
                    ++i;
                    continue;
                }
                if(i < s.size() && s[i] == '}')
                    return ++i, true;
                return fail("expected ',' or '}'");
            }
        }
        if(c == '[')
        {
            v->kind = JsonValue::Array;
            ++i;
            ws();
            if(i < s.size() && s[i] == ']')
                return ++i, true;
            for(;;)
            {
                JsonValue child;
                if(!parseValue(&child, depth + 1))
                    return false;
                v->array.push_back(std::move(child));
                ws();
                if(i < s.size() && s[i] == ',')
                {
                    ++i;
                    continue;
                }
                if(i < s.size() && s[i] == ']')
                    return ++i, true;
                return fail("expected ',' or ']'");
            }
        }
        if(c == '"')
        {
            v->kind = JsonValue::String;
            return parseString(&v->string);
        }
        if(s.compare(i, 4, "true") == 0)
            return v->kind = JsonValue::Bool, v->boolean = true, i += 4, true;
        if(s.compare(i, 5, "false") == 0)
            return v->kind = JsonValue::Bool, v->boolean = false, i += 5, true;
        if(s.compare(i, 4, "null") == 0)
            return v->kind = JsonValue::Null, i += 4, true;
        char* end = nullptr;
        const double d = std::strtod(s.c_str() + i, &end);
        if(end == s.c_str() + i)
            return fail("unexpected character");
        v->kind = JsonValue::Number;
        v->number = d;
        i = (size_t)(end - s.c_str());
        return true;
    }
};

bool numberField(const JsonValue& obj, const char* key, double lo, double hi, double* out, std::string* error)
{
    const JsonValue* v = obj.find(key);
    if(!v)
        return true;
    if(v->kind != JsonValue::Number || !(v->number >= lo && v->number <= hi))
    {
        *error = std::string("parameter '") + key + "' is not a number in range";
        return false;
    }
    *out = v->number;
    return true;
}
} // namespace

bool parsePluginConfig(const std::string& jsonText, std::vector<PluginEntry>* out, std::string* error)
{
    std::string dummy;
    std::string& err = error ? *error : dummy;
    out->clear();
    JsonParser p(jsonText);
    JsonValue root;
    if(!p.parseValue(&root))
        return err = p.error, false;
    p.ws();
    if(p.i != jsonText.size())
        return err = "trailing characters after the JSON document", false;
    const JsonValue* plugins = root.kind == JsonValue::Object ? root.find("plugins") : nullptr;
    if(!plugins || plugins->kind != JsonValue::Array)
        return err = "top-level object must hold a \"plugins\" array", false;
    for(const JsonValue& e : plugins->array)
    {
        if(e.kind != JsonValue::Object)
            return err = "plugins[] entries must be objects", false;
        PluginEntry entry;
        if(const JsonValue* lib = e.find("library"))
        {
            if(lib->kind != JsonValue::String)
                return err = "\"library\" must be a string", false;
            entry.library = lib->string;
        }
        if(const JsonValue* fm = e.find("factoryMethods"))
        {
            if(fm->kind != JsonValue::Object)
                return err = "\"factoryMethods\" must be an object", false;
            for(const auto& kv : fm->object)
            {
                if(kv.second.kind != JsonValue::String)
                    return err = "factoryMethods values must be strings", false;
                if(kv.first == "featureDetector")
                    entry.detectorFactory = kv.second.string;
                else if(kv.first == "featureMatcher")
                    entry.matcherFactory = kv.second.string;
                // other roles of the reference's fixture (loopDetection, loopClosure) are not ours: ignored
            }
        }
        if(const JsonValue* prm = e.find("parameters"))
        {
            if(prm->kind != JsonValue::Object)
                return err = "\"parameters\" must be an object", false;
            OrbB200Options& o = entry.options;
            double v;
            v = o.device;        if(!numberField(*prm, "device", 0, 1023, &v, &err)) return false;        o.device = (int)v;
            v = o.maxWidth;      if(!numberField(*prm, "maxWidth", 1, 4095, &v, &err)) return false;      o.maxWidth = (int)v;
            v = o.maxHeight;     if(!numberField(*prm, "maxHeight", 1, 4095, &v, &err)) return false;     o.maxHeight = (int)v;
            v = o.nfeatures;     if(!numberField(*prm, "nfeatures", 1, 65535, &v, &err)) return false;    o.nfeatures = (int)v;
            v = o.maxBatch;      if(!numberField(*prm, "maxBatch", 1, 4096, &v, &err)) return false;      o.maxBatch = (int)v;
            v = o.nlevels;       if(!numberField(*prm, "nlevels", 1, MSLAM_B200_MAX_LEVELS, &v, &err)) return false; o.nlevels = (int)v;
            v = o.scaleFactor;   if(!numberField(*prm, "scaleFactor", 1.0001, 4.0, &v, &err)) return false; o.scaleFactor = (float)v;
            v = o.fastThreshold; if(!numberField(*prm, "fastThreshold", 1, 254, &v, &err)) return false;  o.fastThreshold = (int)v;
            v = o.ratio;         if(!numberField(*prm, "ratio", 1e-6, 1e6, &v, &err)) return false;       o.ratio = v;
            v = o.maxDescriptors; if(!numberField(*prm, "maxDescriptors", 0, 65535, &v, &err)) return false; o.maxDescriptors = (int)v;
            v = o.iniFastThreshold; if(!numberField(*prm, "iniFastThreshold", 1, 254, &v, &err)) return false; o.iniFastThreshold = (int)v;
            v = o.minFastThreshold; if(!numberField(*prm, "minFastThreshold", 1, 254, &v, &err)) return false; o.minFastThreshold = (int)v;
            v = o.minSize;       if(!numberField(*prm, "minSize", 1, 1e9, &v, &err)) return false;         o.minSize = (int)v;
        }
        out->push_back(entry);
    }
    return true;
}

} // namespace mslam

// the aliases loadFactoryMethod<IOrbFeatureDetector>(lib, "orbDetectorFactory") / <IOrbMatcher>(lib, "orbMatcherFactory") import
MSLAM_B200_DLL_ALIAS(mslam::createOrbB200Detector, orbDetectorFactory)
MSLAM_B200_DLL_ALIAS(mslam::createOrbB200Matcher, orbMatcherFactory)
// the detector the reference's Viewer wires in (viewer.cpp:183), as a second alias of the same interface
MSLAM_B200_DLL_ALIAS(mslam::createDistributedOrbB200Detector, distributedOrbDetectorFactory)
MSLAM_B200_DLL_ALIAS(mslam::createRansacPnpB200, ransacPnpFactory)
