// test_update_text.cpp -- the reference's text-format unit tests replayed against
// oryx_b200/host/update_stream.cpp.  CPU only (no device call).  Paths relative to the reference root:
//   TUT = framework/oryx-common/src/test/java/com/cloudera/oryx/common/text/TextUtilsTest.java
//   MLT = app/oryx-app-common/src/test/java/com/cloudera/oryx/app/common/fn/MLFunctionsTest.java
//   DOC = src/site/markdown/docs/endusers.md
//   GEN = app/oryx-app/src/test/java/com/cloudera/oryx/app/speed/als/MockALSModelUpdateGenerator.java
#include <cmath>
#include <cstdio>
#include <string>
#include <vector>

#include "update_stream.hpp"

using namespace oryx::host;
using V = std::vector<std::string>;

static int g_checks = 0, g_fail = 0;
#define CHECK(cond)                                                                  \
  do {                                                                               \
    g_checks++;                                                                      \
    if (!(cond)) { g_fail++; std::printf("FAIL %s:%d: %s\n", __FILE__, __LINE__, #cond); } \
  } while (0)
template <typename Fn>
static bool throwsBad(Fn fn) {
  try { fn(); } catch (const BadMessage&) { return true; } catch (...) { return false; }
  return false;
}

int main() {
  // TUT:35-42 testParseJSON (and MLT:27-37)
  CHECK((TextUtils::parseJSONArray("[\"a\",\"1\",\"foo\"]") == V{"a", "1", "foo"}));
  CHECK((TextUtils::parseJSONArray("[\"a\",\"1\",\"foo\",\"\"]") == V{"a", "1", "foo", ""}));
  CHECK((TextUtils::parseJSONArray("[\"2.3\"]") == V{"2.3"}));
  CHECK((TextUtils::parseJSONArray("[]") == V{}));
  // TUT:45-66 testParseDelimited
  CHECK((TextUtils::parseDelimited("a,1,foo", ',') == V{"a", "1", "foo"}));
  CHECK((TextUtils::parseDelimited("a,1,foo,", ',') == V{"a", "1", "foo", ""}));
  CHECK((TextUtils::parseDelimited("2.3", ',') == V{"2.3"}));
  CHECK((TextUtils::parseDelimited("\"\"\"a\"\"\"", ',') == V{"\"a\""}));
  CHECK((TextUtils::parseDelimited("\"\"\"\" \"\"\"\"\"\"", ' ') == V{"\"", "\"\""}));
  CHECK((TextUtils::parseDelimited("", ',') == V{""}));
  CHECK((TextUtils::parseDelimited("a\t1,\t,foo", '\t') == V{"a", "1,", ",foo"}));
  CHECK((TextUtils::parseDelimited("a 1 foo ", ' ') == V{"a", "1", "foo", ""}));
  CHECK((TextUtils::parseDelimited("-1.0 a\"\\ \"b", ' ') == V{"-1.0", "a\" \"b"}));
  CHECK((TextUtils::parseDelimited("-1.0 \"a\\\"b\\\"c\"", ' ') == V{"-1.0", "a\"b\"c"}));
  // TUT:69-78 testParsePMMLDelimited
  CHECK((TextUtils::parsePMMLDelimited("1 22 3") == V{"1", "22", "3"}));
  CHECK((TextUtils::parsePMMLDelimited("ab  \"a b\"   \"with \\\"quotes\\\" \" ") == V{"ab", "a b", "with \"quotes\" "}));
  CHECK((TextUtils::parsePMMLDelimited("\"\\\" \\\"\"") == V{"\" \""}));
  CHECK((TextUtils::parsePMMLDelimited(" \" c\\\" d \\\"e \" \" c\\\" d \\\"e \" ") == V{" c\" d \"e ", " c\" d \"e "}));
  // TUT:107-111 testJoinJSON
  CHECK(TextUtils::joinJSON({"1", "2", "3"}) == "[\"1\",\"2\",\"3\"]");
  CHECK(TextUtils::joinJSON({"1 ", "2 ", "3"}) == "[\"1 \",\"2 \",\"3\"]");
  CHECK(TextUtils::joinJSON({}) == "[]");
  // TUT:131-135 testReadJSON: "[1,2]" as float[] -> {1.0f, 2.0f}; ["foo", "bar"] as a List
  {
    std::string id;
    std::vector<float> v;
    parseFeatureMessage("[\"k\",[1,2]]", false, nullptr, &id, &v, nullptr, nullptr);
    CHECK(id == "k" && v.size() == 2 && v[0] == 1.0f && v[1] == 2.0f);
    CHECK((TextUtils::parseJSONArray("[\"foo\", \"bar\"]") == V{"foo", "bar"}));
  }
  // MLT:39-50 testParseCSV
  CHECK((MLFunctions::parse("a,1,foo") == V{"a", "1", "foo"}));
  CHECK((MLFunctions::parse("a,1,foo,") == V{"a", "1", "foo", ""}));
  CHECK((MLFunctions::parse("2.3") == V{"2.3"}));
  CHECK((MLFunctions::parse("") == V{""}));
  // MLT:52-58 testToTimestamp
  CHECK(MLFunctions::toTimestamp("a,b,c,123") == 123);
  CHECK(MLFunctions::toTimestamp("a,b,c,123,") == 123);
  CHECK(MLFunctions::toTimestamp("[\"a\",\"b\",\"c\",123]") == 123);
  CHECK(MLFunctions::toTimestamp("[\"a\",\"b\",\"c\",123,\"d\"]") == 123);
  // MLT:60-67 testSumWithNaN
  const double NaN = std::nan("");
  CHECK(MLFunctions::sumWithNaN({1.0}) == 1.0);
  CHECK(MLFunctions::sumWithNaN({1.0, 2.0, 3.0}) == 6.0);
  CHECK(MLFunctions::sumWithNaN({1.0, NaN, 3.0}) == 3.0);
  CHECK(std::isnan(MLFunctions::sumWithNaN({1.0, 2.0, NaN})));
  CHECK(std::isnan(MLFunctions::sumWithNaN({NaN})));
  // MLT:69-72 testParseBadLine (IOException), :74-77 testParseBadTimestamp (ArrayIndexOutOfBounds)
  CHECK(throwsBad([] { MLFunctions::parse("[1,]"); }));
  {
    bool oob = false;
    try { MLFunctions::toTimestamp("[1,2,3]"); } catch (const std::out_of_range&) { oob = true; }
    CHECK(oob);
  }

  // DOC:103-114: the model.pmml example (completed into a well-formed document)
  {
    const char* pmml =
        "<?xml version=\"1.0\" encoding=\"UTF-8\" standalone=\"yes\"?>\n"
        "<PMML xmlns=\"http://www.dmg.org/PMML-4_2\" version=\"4.2.1\">\n"
        "    <Header>\n        <Application name=\"Oryx\"/>\n        <Timestamp>2014-12-18T04:48:54-0800</Timestamp>\n    </Header>\n"
        "    <Extension name=\"X\" value=\"X/\"/>\n    <Extension name=\"Y\" value=\"Y/\"/>\n"
        "    <Extension name=\"features\" value=\"10\"/>\n    <Extension name=\"lambda\" value=\"0.001\"/>\n"
        "    <Extension name=\"implicit\" value=\"true\"/>\n    <Extension name=\"alpha\" value=\"1.0\"/>\n"
        "    <Extension name=\"XIDs\">56 168 222 343 397</Extension>\n"
        "    <Extension name=\"YIDs\">ab \"a b\" \"with \\&quot;quotes\\&quot; \"</Extension>\n"
        "</PMML>\n";
    auto ext = PMMLExtensions::fromString(pmml);
    CHECK(ext.values["X"] == "X/" && ext.values["Y"] == "Y/");
    CHECK(ext.values["features"] == "10" && ext.values["implicit"] == "true");
    CHECK(ext.values["lambda"] == "0.001" && ext.values["alpha"] == "1.0");
    CHECK((ext.contents["XIDs"] == V{"56", "168", "222", "343", "397"}));
    CHECK((ext.contents["YIDs"] == V{"ab", "a b", "with \"quotes\" "}));  // PMML quoting, TUT:72-73
    CHECK(!ext.hasValue("XIDs") && !ext.contents.count("nope"));
    CHECK(throwsBad([] { PMMLExtensions::fromString("<NotPMML/>"); }));
  }
  // DOC:120-128: feature part-file lines; numeric key, vector printed from doubles
  {
    std::string id;
    std::vector<float> v;
    parseFeatureMessage("[56,[0.5746282834154238,-0.08896614131333057,-0.029456222765775263,"
                        "0.6039821219690552,0.1497901814774658,-0.018654312114339863,-0.37342063488340266,"
                        "-0.2370768843521807,1.148260034028485,1.0645643656769153]]",
                        false, nullptr, &id, &v, nullptr, nullptr);
    CHECK(id == "56" && v.size() == 10);
    CHECK(v[0] == (float)0.5746282834154238 && v[8] == (float)1.148260034028485 && v[5] == (float)-0.018654312114339863);
  }
  // GEN:67-100: what the generator emits for id 6 (X) and id 1 (Y)
  {
    const float x6[2] = {-0.6790019f, 0.1732324f};
    std::string other = "1";
    std::string msg = toUpdateJSON('X', "6", x6, 2, &other);
    CHECK(msg == "[\"X\",\"6\",[-0.6790019,0.1732324],[\"1\"]]");
    char matrix;
    std::string id;
    std::vector<float> v;
    bool has;
    V others;
    parseFeatureMessage("[\"X\",\"6\",[-0.6790019,0.1732324],[\"1\",\"4\"]]", true, &matrix, &id, &v, &has, &others);
    CHECK(matrix == 'X' && id == "6" && v.size() == 2 && v[0] == x6[0] && v[1] == x6[1] && has && (others == V{"1", "4"}));
    parseFeatureMessage("[\"Y\",\"1\",[-0.7203235,0.45654634]]", true, &matrix, &id, &v, &has, &others);
    CHECK(matrix == 'Y' && id == "1" && v[0] == -0.7203235f && v[1] == 0.45654634f && !has && others.empty());
    CHECK(throwsBad([] { parseFeatureMessage("[\"Z\",\"1\",[1.0]]", true, nullptr, nullptr, nullptr, nullptr, nullptr); }));
    CHECK(throwsBad([] { parseFeatureMessage("[\"X\",\"1\"]", true, nullptr, nullptr, nullptr, nullptr, nullptr); }));
    CHECK(throwsBad([] { parseFeatureMessage("[\"X\",\"1\",[1.0,]]", true, nullptr, nullptr, nullptr, nullptr, nullptr); }));
    CHECK(throwsBad([] { parseFeatureMessage("[\"X\",\"1\",[01]]", true, nullptr, nullptr, nullptr, nullptr, nullptr); }));
  }
  // Float.toString layout (JLS / Float.toString javadoc examples)
  CHECK(TextUtils::floatToString(1.0f) == "1.0" && TextUtils::floatToString(-0.0f) == "-0.0");
  CHECK(TextUtils::floatToString(1.0e-5f) == "1.0E-5" && TextUtils::floatToString(0.001f) == "0.001");
  CHECK(TextUtils::floatToString(1.0e7f) == "1.0E7" && TextUtils::floatToString(9999999.0f) == "9999999.0");
  CHECK(TextUtils::floatToString(3.4028235e38f) == "3.4028235E38" && TextUtils::floatToString(1.4e-45f) == "1.4E-45");
  CHECK(TextUtils::floatToString(0.13043478f) == "0.13043478" && TextUtils::floatToString(100.0f) == "100.0");
  CHECK(TextUtils::doubleToString(0.5746282834154238) == "0.5746282834154238" && TextUtils::doubleToString(4.9e-324) == "4.9E-324");

  std::printf("%d checks, %d failed\n", g_checks, g_fail);
  return g_fail ? 1 : 0;
}
