// TEST INFRASTRUCTURE ONLY: the handful of PS::CMeshVV members the reference's scalar polygonizer calls
// (Parsip100/ParsipHaptics/include/CPolyParsipOptimized.cpp:19-20, 300-316, 432-441, 504-505, 594-613), written against the
// reference's own class declaration (PS_FrameWork/include/PS_MeshVV.h). The reference's PS_MeshVV.cpp cannot be compiled here:
// it includes GL/glew.h for its draw calls (no graphics stack on this box). These are plain vector appends, no arithmetic.
#include "PS_FrameWork/include/PS_MeshVV.h"
namespace PS {
CMeshVV::CMeshVV(POLYGON_MODE mode, int vertexSize, int colorSize, int texChannels, int texCoordSize) {
    initMesh(mode, vertexSize, colorSize, texChannels, texCoordSize);
}
CMeshVV::CMeshVV(const CMeshVV& other) { copyFrom(other); }
CMeshVV::~CMeshVV(void) { removeAll(); }
void CMeshVV::initMesh(POLYGON_MODE mode, int szUnitVertex, int szUnitColor, int texChannels, int szUnitTexCoord) {
    m_faceMode = mode;
    m_szUnitVertex = szUnitVertex;
    m_szUnitColor = szUnitColor;
    m_szTexChannels = texChannels;
    for (int i = 0; i < MAX_TEXTURE_CHANNELS; i++) m_szUnitTexCoord[i] = szUnitTexCoord;
}
void CMeshVV::removeAll(void) {
    m_lstLabels.resize(0); m_lstColors.resize(0); m_lstVertices.resize(0); m_lstNormals.resize(0); m_lstFaces.resize(0);
    for (int i = 0; i < MAX_TEXTURE_CHANNELS; i++) m_lstTexChannels[i].resize(0);
}
void CMeshVV::addVertex(float x, float y, float z) { m_lstVertices.push_back(x); m_lstVertices.push_back(y); m_lstVertices.push_back(z); }
void CMeshVV::addNormal(float x, float y, float z) { m_lstNormals.push_back(x); m_lstNormals.push_back(y); m_lstNormals.push_back(z); }
void CMeshVV::addColor(vec4f c) {
    const float v[4] = {c.x, c.y, c.z, c.w};
    for (int i = 0; i < m_szUnitColor && i < 4; i++) m_lstColors.push_back(v[i]);
}
void CMeshVV::addTriangle(U32 a, U32 b, U32 c) { m_lstFaces.push_back(a); m_lstFaces.push_back(b); m_lstFaces.push_back(c); }
int CMeshVV::getUnitFaceSize(void) const { return m_faceMode == TRIANGLES ? 3 : (m_faceMode == QUADS ? 4 : (m_faceMode == LINES ? 2 : 1)); }
U32 CMeshVV::countFaces(void) const { return (U32)(m_lstFaces.size() / (size_t)getUnitFaceSize()); }
U32 CMeshVV::countVertices(void) const { return m_szUnitVertex > 0 ? (U32)(m_lstVertices.size() / (size_t)m_szUnitVertex) : 0; }
bool CMeshVV::copyFrom(const CMeshVV& rhs) {
    m_faceMode = rhs.m_faceMode; m_szUnitVertex = rhs.m_szUnitVertex; m_szUnitColor = rhs.m_szUnitColor; m_szTexChannels = rhs.m_szTexChannels;
    m_lstLabels = rhs.m_lstLabels; m_lstColors = rhs.m_lstColors; m_lstVertices = rhs.m_lstVertices; m_lstNormals = rhs.m_lstNormals;
    m_lstFaces = rhs.m_lstFaces;
    for (int i = 0; i < MAX_TEXTURE_CHANNELS; i++) { m_szUnitTexCoord[i] = rhs.m_szUnitTexCoord[i]; m_lstTexChannels[i] = rhs.m_lstTexChannels[i]; }
    return true;
}
bool CMeshVV::appendFrom(const CMeshVV& rhs) {
    if (m_faceMode != rhs.m_faceMode || m_szUnitVertex != rhs.m_szUnitVertex || m_szUnitColor != rhs.m_szUnitColor) return false;
    const U32 offset = countVertices();
    m_lstLabels.insert(m_lstLabels.end(), rhs.m_lstLabels.begin(), rhs.m_lstLabels.end());
    m_lstColors.insert(m_lstColors.end(), rhs.m_lstColors.begin(), rhs.m_lstColors.end());
    m_lstNormals.insert(m_lstNormals.end(), rhs.m_lstNormals.begin(), rhs.m_lstNormals.end());
    m_lstVertices.insert(m_lstVertices.end(), rhs.m_lstVertices.begin(), rhs.m_lstVertices.end());
    for (size_t i = 0; i < rhs.m_lstFaces.size(); i++) m_lstFaces.push_back(rhs.m_lstFaces[i] + offset);
    return true;
}
void CMeshVV::drawBuffered() const {}
void CMeshVV::drawNormals(int) const {}
}
