#include "vh_graph.h"
#include <algorithm>
#include <cstring>

namespace vh {

// ---- passes: reference src/render_core/pass/{integrate,accumulate,tonemapping,gamma}.cpp ----------------------------------------
class IntegratePass : public RenderPass {            // radiance = (Li, 1) of one frame index
public:
    explicit IntegratePass(const PassDesc &d) : RenderPass(d) {}
    std::string impl_type() const override { return "integrate"; }
    ChannelList outputs() const override { return {{"radiance", "radiance", false}}; }
    int output_plane(const std::string &) const override { return VB_FILM_RT; }
    int input_plane(const std::string &) const override { return -1; }
    bool feeds_accumulate = false;                    // set by CustomizedPipeline after set-up
    void dispatch(Pipeline &p, uint32_t frame_index) override {
        // one launch integrates the frame into RT and folds it into ACCUM (lerp(1 / (k + 1), ACCUM, RT), k = frames accumulated so far = the
        // frame index here): the accumulate pass that follows is fused into this call, its own dispatch has nothing left to do
        if (!feeds_accumulate) throw std::runtime_error("render graph: the integrate pass must feed an accumulate pass (fused launch)");
        p.check(vb_render_frames(p.ctx, frame_index, 1, 1, VB_ACCUM_LERP), "vb_render_frames");
    }
};
VH_MAKE_CLASS_CREATOR(IntegratePass, pass, integrate, constructor)

class AccumulatePass : public RenderPass {           // output = lerp(1 / (frame_index + 1), output, input), w = 1
public:
    explicit AccumulatePass(const PassDesc &d) : RenderPass(d) {}
    std::string impl_type() const override { return "accumulate"; }
    ChannelList inputs() const override { return {{"input", "input", false}}; }
    ChannelList outputs() const override { return {{"output", "output", false}}; }
    int output_plane(const std::string &) const override { return VB_FILM_ACCUM; }
    int input_plane(const std::string &) const override { return VB_FILM_RT; }
    void dispatch(Pipeline &, uint32_t) override {}   // done by the integrate pass's launch (see there)
};
VH_MAKE_CLASS_CREATOR(AccumulatePass, pass, accumulate, constructor)

class ToneMappingPass : public RenderPass {          // output = (tone_mapper(input.rgb), 1); the pass owns a default ("linear") tone mapper
public:
    std::shared_ptr<ToneMapper> tone_mapper;
    explicit ToneMappingPass(const PassDesc &d) : RenderPass(d) {
        ToneMapperDesc td;
        Json empty; empty.kind = Json::Object;
        td.init(ParameterSet(empty));                 // tonemapping.cpp:19-23: ToneMapperDesc::init(DataWrap::object()) -> "linear"
        tone_mapper = create_shared<ToneMapper>(td);
    }
    std::string impl_type() const override { return "tonemapping"; }
    ChannelList inputs() const override { return {{"input", "input", false}}; }
    ChannelList outputs() const override { return {{"output", "output", false}}; }
    int output_plane(const std::string &) const override { return VB_FILM_OUTPUT; }
    int input_plane(const std::string &) const override { return VB_FILM_ACCUM; }
    void dispatch(Pipeline &p, uint32_t) override { p.check(vb_pass_tonemapping(p.ctx, tone_mapper->device_tag()), "vb_pass_tonemapping"); }
};
VH_MAKE_CLASS_CREATOR(ToneMappingPass, pass, tonemapping, constructor)

class GammaCorrectionPass : public RenderPass {      // output = linear_to_srgb(input), all four channels
public:
    explicit GammaCorrectionPass(const PassDesc &d) : RenderPass(d) {}
    std::string impl_type() const override { return "gamma"; }
    ChannelList inputs() const override { return {{"input", "input", false}}; }
    ChannelList outputs() const override { return {{"output", "output", false}}; }
    int output_plane(const std::string &) const override { return VB_FILM_VIEW; }
    int input_plane(const std::string &) const override { return VB_FILM_OUTPUT; }
    void dispatch(Pipeline &p, uint32_t) override { p.check(vb_pass_gamma(p.ctx), "vb_pass_gamma"); }
};
VH_MAKE_CLASS_CREATOR(GammaCorrectionPass, pass, gamma, constructor)

// ---- graph: reference src/render_graph/graph.cpp:37-104 -------------------------------------------------------------------------
RenderPass *RenderGraph::pass_at(const std::string &name) const {
    auto it = pass_map_.find(name);
    if (it == pass_map_.end()) throw std::runtime_error("render graph: no pass named '" + name + "'");
    return it->second.get();
}
RenderGraph::EdgeData RenderGraph::find_edge(const Field &dst) const {
    for (const auto &e : edges_) if (dst.str() == e.dst.str()) return e;
    return {};
}
void RenderGraph::dfs(RenderPass *pass, std::vector<RenderPass *> &on_path) {
    if (std::find(on_path.begin(), on_path.end(), pass) != on_path.end()) throw std::runtime_error("render graph: cycle through pass '" + pass->pass_name + "'");
    on_path.push_back(pass);
    for (const auto &in : pass->inputs()) {
        EdgeData e = find_edge(Field(pass->pass_name + "." + in.name));
        if (e.empty()) {
            if (!in.optional) throw std::runtime_error("render graph: input '" + pass->pass_name + "." + in.name + "' is not connected");
            continue;
        }
        RenderPass *src = pass_at(e.src.pass);
        bool has = false;
        for (const auto &o : src->outputs()) has |= o.name == e.src.channel;
        if (!has) throw std::runtime_error("render graph: pass '" + e.src.pass + "' has no output '" + e.src.channel + "'");
        simple_edges_.push_back(e);
        dfs(src, on_path);
    }
    on_path.pop_back();
    if (std::find(pass_list_.begin(), pass_list_.end(), pass) == pass_list_.end()) pass_list_.push_back(pass);      // PassList::push_back: once
}
void RenderGraph::setup() {
    if (output_.empty()) throw std::runtime_error("render graph: no output marked");
    pass_list_.clear(); simple_edges_.clear();
    RenderPass *out = pass_at(output_.pass);
    bool has = false;
    for (const auto &o : out->outputs()) has |= o.name == output_.channel;
    if (!has) throw std::runtime_error("render graph: pass '" + output_.pass + "' has no output '" + output_.channel + "'");
    std::vector<RenderPass *> on_path;
    dfs(out, on_path);                                            // _simplification
    for (const auto &e : simple_edges_) {                         // _allocate_resource: an edge = the producer's film plane
        RenderPass *src = pass_at(e.src.pass), *dst = pass_at(e.dst.pass);
        const int plane = src->output_plane(e.src.channel);
        if (dst->input_plane(e.dst.channel) != plane)
            throw std::runtime_error("render graph: edge " + e.src.str() + " -> " + e.dst.str() + " is not one this library's pass kernels can be wired for");
        src->res_map[e.src.channel] = plane; dst->res_map[e.dst.channel] = plane;
    }
    out->res_map[output_.channel] = out->output_plane(output_.channel);
}
int RenderGraph::output_plane() const { return pass_at(output_.pass)->res_map.at(output_.channel); }
void RenderGraph::dispatch(Pipeline &p, uint32_t frame_index) {
    for (RenderPass *pass : pass_list_) pass->dispatch(p, frame_index);
}

// ---- pipeline: reference src/pipelines/customized/pipeline.cpp ------------------------------------------------------------------
void CustomizedPipeline::prepare_render_graph() {
    render_graph.add_pass(RenderPass::create("integrate"), "integrate");
    render_graph.add_pass(RenderPass::create("accumulate"), "accumulate");
    render_graph.add_pass(RenderPass::create("tonemapping"), "tonemapping");
    render_graph.add_pass(RenderPass::create("gamma"), "gamma");
    render_graph.add_edge("integrate.radiance", "accumulate.input");
    render_graph.add_edge("accumulate.output", "tonemapping.input");
    render_graph.add_edge("tonemapping.output", "gamma.input");
    render_graph.mark_output("gamma.output");
    render_graph.setup();
    const auto &pl = render_graph.pass_list();
    for (size_t i = 0; i < pl.size(); ++i)
        if (auto *ip = dynamic_cast<IntegratePass *>(pl[i]))
            ip->feeds_accumulate = i + 1 < pl.size() && dynamic_cast<AccumulatePass *>(pl[i + 1]) != nullptr;
}
void CustomizedPipeline::prepare(int device) {
    if (dist_world > 1) throw std::runtime_error("customized pipeline: frame partitioning over several GPUs is built for the fixed pipeline only");
    Pipeline::prepare(device);
    prepare_render_graph();
}
void CustomizedPipeline::render(uint32_t n_frames) {
    for (uint32_t i = 0; i < n_frames; ++i) {
        render_graph.dispatch(*this, scene.integrator->frame_index);
        scene.integrator->frame_index++;
        after_render();
    }
}
void CustomizedPipeline::final_picture(std::vector<float> &rgba) {
    rgba.resize((size_t)records.width * records.height * 4);
    check(vb_film_download(ctx, rgba.data(), render_graph.output_plane()), "vb_film_download");
}

VH_MAKE_CLASS_CREATOR(CustomizedPipeline, pipeline, customized, constructor)

}  // namespace vh
