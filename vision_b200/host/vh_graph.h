// Render graph of the `customized` pipeline: passes with named input / output channels, edges "pass.channel" -> "pass.channel", one
// marked output; set-up walks back from the output pass (unreached passes are dropped), orders the passes and binds every edge to a
// buffer.   reference src/render_graph/graph.{h,cpp}, pass.h; src/render_core/pass/*.cpp; src/pipelines/customized/pipeline.cpp
//
// Buffers: the reference allocates one float4 buffer per edge; here an edge is bound to the film plane of the C ABI that the producing
// pass's kernel writes (integrate.radiance -> VB_FILM_RT, accumulate.output -> VB_FILM_ACCUM, tonemapping.output -> VB_FILM_OUTPUT,
// gamma.output -> VB_FILM_VIEW), and set-up checks that each consumer reads the plane its kernel reads - a graph that wires the
// four passes differently is refused by name, never run differently.
#pragma once
#include "vh_scene.h"
#include <list>
#include <map>

namespace vh {

struct ChannelDesc { std::string name, desc; bool optional; };
using ChannelList = std::vector<ChannelDesc>;

class RenderPass : public Node {
public:
    using Desc = PassDesc;
    explicit RenderPass(const PassDesc &) {}
    std::map<std::string, int> res_map;                          // channel -> film plane (VB_FILM_*)
    std::string pass_name;                                       // name inside the graph
    virtual ChannelList inputs() const { return {}; }
    virtual ChannelList outputs() const { return {}; }
    virtual int output_plane(const std::string &channel) const = 0;     // plane this pass's kernel writes for `channel`
    virtual int input_plane(const std::string &channel) const = 0;      // plane this pass's kernel reads for `channel`
    virtual void dispatch(Pipeline &p, uint32_t frame_index) = 0;
    static std::shared_ptr<RenderPass> create(const std::string &name) { return create_shared<RenderPass>(PassDesc(name)); }
};

class RenderGraph {
public:
    struct Field {
        std::string pass, channel;
        Field() = default;
        explicit Field(const std::string &full) { size_t dot = full.find_last_of('.'); pass = full.substr(0, dot); channel = dot == std::string::npos ? "" : full.substr(dot + 1); }
        std::string str() const { return pass + "." + channel; }
        bool empty() const { return pass.empty() || channel.empty(); }
    };
    struct EdgeData { Field src, dst; bool empty() const { return src.empty() || dst.empty(); } };

    void add_pass(std::shared_ptr<RenderPass> pass, const std::string &name) { pass->pass_name = name; pass_map_[name] = std::move(pass); }
    void add_edge(const std::string &output, const std::string &input) { edges_.push_back(EdgeData{Field(output), Field(input)}); }
    void mark_output(const std::string &output) { output_ = Field(output); }
    void setup();                                                // simplification (DFS from the output pass) + resource binding
    void dispatch(Pipeline &p, uint32_t frame_index);            // the ordered passes of one frame
    int output_plane() const;                                    // film plane of the marked output
    const std::vector<RenderPass *> &pass_list() const { return pass_list_; }

private:
    std::map<std::string, std::shared_ptr<RenderPass>> pass_map_;
    std::list<EdgeData> edges_, simple_edges_;
    Field output_;
    std::vector<RenderPass *> pass_list_;
    EdgeData find_edge(const Field &dst) const;
    void dfs(RenderPass *pass, std::vector<RenderPass *> &on_path);
    RenderPass *pass_at(const std::string &name) const;
};

class CustomizedPipeline : public Pipeline {     // "vision-pipeline-customized"
public:
    explicit CustomizedPipeline(const PipelineDesc &d) : Pipeline(d) {}
    std::string impl_type() const override { return "customized"; }
    RenderGraph render_graph;
    void prepare_render_graph();                                 // customized/pipeline.cpp:28-44
    void prepare(int device) override;
    void render(uint32_t n_frames) override;                     // per frame: the graph's passes, then after_render()
    void final_picture(std::vector<float> &rgba) override;       // the graph's output buffer (already gamma-corrected)
    bool output_is_display_referred() const override { return true; }
};

}  // namespace vh
