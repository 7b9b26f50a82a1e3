/*
 * ref_stubs.cpp — TEST INFRASTRUCTURE ONLY (part of oracle/_ref/libsparkle_ref.so).
 *
 * Link-time stand-ins for reference symbols that live in translation units which cannot be built here (RHI back ends,
 * UI, cvar registry, texture codecs) and that the CPU path tracer never calls on the path the harness drives
 * (CPURenderer::BasePass/DenoisePass/ToneMappingPass, SceneRenderProxy::Intersect). Every stub that would indicate a
 * wrong turn aborts loudly. No reference code is reproduced here.
 */
#include "core/FileManager.h"
#include "core/Path.h"
#include "core/ThreadUtils.h"
#include "io/Image.h"
#include "io/TextureCompression.h"
#include "renderer/BindlessManager.h"
#include "renderer/RenderConfig.h"
#include "renderer/pass/ScreenQuadPass.h"
#include "renderer/pass/UiPass.h"
#include "renderer/renderer/Renderer.h"
#include "core/cook/Cooker.h"
#include "renderer/pass/IBLPass.h"
#include "renderer/resource/ImageBasedLighting.h"
#include "rhi/RHI.h"
#include "application/UiManager.h"

#include <imgui.h>

#include <cstdio>
#include <cstdlib>

#define SPK_STUB_ABORT()                                                                                               \
    do                                                                                                                 \
    {                                                                                                                  \
        std::fprintf(stderr, "oracle/_ref: stub %s reached — not on the CPU path-tracer path\n", __func__);           \
        std::abort();                                                                                                  \
    } while (0)

namespace sparkle
{
// ---- Renderer base (Renderer.cpp pulls in the GPU/raster renderers; only the ctor is needed) ----
Renderer::Renderer(const RenderConfig &render_config, RHIContext *rhi_context, SceneRenderProxy *scene_render_proxy)
    : rhi_(rhi_context), scene_render_proxy_(scene_render_proxy), resolution_(render_config.GetResolution()),
      render_config_(render_config)
{
}
bool Renderer::IsReadyForAutoScreenshot() const
{
    return scene_loaded_;
}
bool Renderer::ReadbackFinalOutputIfRequested(RHIRenderTarget *, bool, RHIPipelineStage)
{
    SPK_STUB_ABORT();
}

void RenderConfig::Validate()
{
}

// ---- passes only reached from CPURenderer::InitRenderResources/Render (RHI presentation) ----
void ScreenQuadPass::Render() { SPK_STUB_ABORT(); }
void ScreenQuadPass::InitRenderResources(const RenderConfig &) { SPK_STUB_ABORT(); }
void ScreenQuadPass::UpdateFrameData(const RenderConfig &, SceneRenderProxy *) { SPK_STUB_ABORT(); }
void ScreenQuadPass::SetupRenderPass() { SPK_STUB_ABORT(); }
void ScreenQuadPass::SetupPipeline() { SPK_STUB_ABORT(); }
void ScreenQuadPass::SetupVertices() { SPK_STUB_ABORT(); }
void ScreenQuadPass::SetupVertexShader() { SPK_STUB_ABORT(); }
void ScreenQuadPass::SetupPixelShader() { SPK_STUB_ABORT(); }
void ScreenQuadPass::CompilePipeline() { SPK_STUB_ABORT(); }
void ScreenQuadPass::BindVertexShaderResources() { SPK_STUB_ABORT(); }
void ScreenQuadPass::BindPixelShaderResources() { SPK_STUB_ABORT(); }
void UiPass::Render() { SPK_STUB_ABORT(); }
void UiPass::InitRenderResources(const RenderConfig &) { SPK_STUB_ABORT(); }

// ---- RHI / bindless / IBL ----
void BindlessManager::InitRenderResources(RHIContext *) { SPK_STUB_ABORT(); }
void BindlessManager::UpdateFrameData(RHIContext *) { SPK_STUB_ABORT(); }
void BindlessManager::RegisterTexture(const RHIResourceRef<RHIImage> &) { SPK_STUB_ABORT(); }
void BindlessManager::UnregisterTexture(RHIImage *) { SPK_STUB_ABORT(); }
void RHIBuffer::Upload(RHIContext *, const void *) { SPK_STUB_ABORT(); }
void RHIBuffer::UploadImmediate(const void *) { SPK_STUB_ABORT(); }
RHIResourceRef<RHIImage> RHIContext::CreateTexture(const Image2D *, const std::string &) { SPK_STUB_ABORT(); }
RHIResourceRef<RHIImage> RHIContext::CreateTextureCube(const Image2DCube *, const std::string &) { SPK_STUB_ABORT(); }
RHIResourceRef<RHIImage> RHIContext::GetOrCreateDummyTexture(RHIImage::Attribute) { SPK_STUB_ABORT(); }
bool RHIContext::IsHeadless() const { return true; }
ImageBasedLighting::ImageBasedLighting(const RHIResourceRef<RHIImage> &, std::shared_ptr<const Image2DCube>)
{
    SPK_STUB_ABORT();
}
ImageBasedLighting::~ImageBasedLighting() = default;
void ImageBasedLighting::InitRenderResources(RHIContext *, const RenderConfig &) { SPK_STUB_ABORT(); }
bool ImageBasedLighting::NeedUpdate() const { return false; }
void ImageBasedLighting::CookOnTheFly(const RenderConfig &) { SPK_STUB_ABORT(); }

// ---- misc ----
FileManager *FileManager::native_file_manager_ = nullptr;
void SetCurrentThreadName(const std::string &) {}
void UiManager::RequestWindowDraw(UiManager::CustomUiWindow &&) { SPK_STUB_ABORT(); }
Image2D TextureCompression::Decode(const Image2D &, unsigned) { SPK_STUB_ABORT(); } // textures are uncompressed RGBA here
} // namespace sparkle

// ---- Dear ImGui entry points referenced by Logger::DrawUi (on-screen log window; never drawn headless) ----
namespace ImGui
{
float GetFontSize() { SPK_STUB_ABORT(); }
ImGuiViewport *GetMainViewport() { SPK_STUB_ABORT(); }
void SetNextWindowPos(const ImVec2 &, ImGuiCond, const ImVec2 &) { SPK_STUB_ABORT(); }
void SetNextWindowSize(const ImVec2 &, ImGuiCond) { SPK_STUB_ABORT(); }
bool Begin(const char *, bool *, ImGuiWindowFlags) { SPK_STUB_ABORT(); }
void End() { SPK_STUB_ABORT(); }
float GetWindowWidth() { SPK_STUB_ABORT(); }
ImVec2 CalcTextSize(const char *, const char *, bool, float) { SPK_STUB_ABORT(); }
void SetCursorPosX(float) { SPK_STUB_ABORT(); }
void TextUnformatted(const char *, const char *) { SPK_STUB_ABORT(); }
} // namespace ImGui
