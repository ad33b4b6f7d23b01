/* TEST INFRASTRUCTURE — command implementations of the oracle harness (included by
 * mts_oracle.cpp). Every numeric result below is produced by the reference's own classes. */

static int cmdTraceImpl(int argc, char **argv) {
	if (argc < 6) { std::cerr << "trace SCENE RAYS OUT closest|closest_full|any|stats [threads] [repeat]\n"; return 1; }
	sio::Scene sf; sio::load(argv[2], sf);
	std::vector<float> rays = sio::loadFloats(argv[3]);
	const size_t n = rays.size() / 8;
	std::string modeStr = argv[5];
	int mode = modeStr == "closest" ? 0 : modeStr == "closest_full" ? 1 : modeStr == "any" ? 2 : modeStr == "stats" ? 3 : -1;
	if (mode < 0) { std::cerr << "bad mode\n"; return 1; }
	int threads = argc > 6 ? atoi(argv[6]) : 1;
	int repeat = argc > 7 ? atoi(argv[7]) : 1;

	Built b = buildScene(sf, "");
	std::vector<float> tuv; std::vector<uint32_t> ids, occl; std::vector<float> full;
	if (mode == 2) occl.resize(n); else { tuv.resize(n * 3); ids.resize(n * 2); }
	if (mode == 1) full.resize(n * 12);
	std::atomic<uint64_t> kdTrav(0), kdIsect(0);
	TraceCtx ctx = { b.scene.get(), rays.data(), mode, tuv.data(), ids.data(), full.data(), occl.data(), &kdTrav, &kdIsect };
	double best = 1e30;
	std::string all;
	for (int r = 0; r < repeat; ++r) {
		double t0 = now();
		parallelFor(n, threads, traceRange, &ctx);
		const double dt = now() - t0;
		best = std::min(best, dt);
		all += (r ? ", " : "") + formatString("%.6f", dt);
	}
	FILE *f = fopen(argv[4], "wb");
	if (!f) { std::cerr << "cannot write output\n"; return 1; }
	if (mode == 2) fwrite(occl.data(), 4, n, f);
	else {
		fwrite(tuv.data(), 4, n * 3, f);
		fwrite(ids.data(), 4, n * 2, f);
		if (mode == 1) fwrite(full.data(), 4, n * 12, f);
	}
	fclose(f);
	size_t hits = 0;
	if (mode == 2) for (size_t i = 0; i < n; ++i) hits += occl[i];
	else if (mode == 3) for (size_t i = 0; i < n; ++i) hits += std::isfinite(tuv[i*3]);
	else for (size_t i = 0; i < n; ++i) hits += ids[i*2] != 0xFFFFFFFFu;
	const AABB &bb = b.scene->getKDTree()->getAABB();
	printf("{\"cmd\": \"trace\", \"mode\": \"%s\", \"rays\": %zu, \"hits\": %zu, \"threads\": %d, "
		"\"kd_build_s\": %.6f, \"trace_s\": %.6f, \"trace_s_all\": [%s], \"mrays_per_s\": %.6f, "
		"\"kd_traversals\": %llu, \"kd_intersections\": %llu, "
		"\"aabb\": [%.9g, %.9g, %.9g, %.9g, %.9g, %.9g]}\n",
		modeStr.c_str(), n, hits, threads, b.buildSeconds, best, all.c_str(), n / best * 1e-6,
		(unsigned long long) (kdTrav.load() / (uint64_t) std::max(repeat, 1)), (unsigned long long) (kdIsect.load() / (uint64_t) std::max(repeat, 1)),
		bb.min.x, bb.min.y, bb.min.z, bb.max.x, bb.max.y, bb.max.z);
	return 0;
}

/* Dump the reference's precomputed TriAccel records (triaccel.h:63-96) and the kd-tree's
 * enlarged AABB (gkdtree.h:1172-1176). */
static int cmdTriAccel(int argc, char **argv) {
	if (argc < 4) { std::cerr << "triaccel SCENE OUT\n"; return 1; }
	sio::Scene sf; sio::load(argv[2], sf);
	Built b = buildScene(sf, "");
	const ShapeKDTree *kd = b.scene->getKDTree();
	size_t n = kd->getPrimitiveCount();
	FILE *f = fopen(argv[3], "wb");
	const AABB &bb = kd->getAABB();
	float box[6] = { bb.min.x, bb.min.y, bb.min.z, bb.max.x, bb.max.y, bb.max.z };
	fwrite(box, 4, 6, f);
	fwrite(Instance::triAccel(kd), sizeof(TriAccel), n, f);
	fclose(f);
	printf("{\"cmd\": \"triaccel\", \"prims\": %zu, \"kd_build_s\": %.6f}\n", n, b.buildSeconds);
	return 0;
}

/* queries: [matId, wi.xyz, wo.xyz, s.x, s.y] -> out 16 floats:
 *   f(wi,wo)[3] pdf(wi,wo)  | 3-arg sample: wo[3] f[3] pdf sampledType | 2-arg sample value [3] | pad */
static int cmdBsdf(int argc, char **argv) {
	if (argc < 5) { std::cerr << "bsdf SCENE QUERIES OUT\n"; return 1; }
	sio::Scene sf; sio::load(argv[2], sf);
	std::vector<float> q = sio::loadFloats(argv[3]);
	const size_t n = q.size() / 9;
	Built b = buildScene(sf, "", false);
	std::vector<float> out(n * 16, 0.0f);
	Intersection its;
	for (size_t i = 0; i < n; ++i) {
		const float *e = &q[i * 9];
		const BSDF *bsdf = b.bsdfs[(size_t) e[0]];
		Vector wi(e[1], e[2], e[3]), wo(e[4], e[5], e[6]);
		Point2 s(e[7], e[8]);
		float *o = &out[i * 16];
		{
			BSDFQueryRecord bRec(its, wi, wo);
			Spectrum f = bsdf->f(bRec);
			f.toLinearRGB(o[0], o[1], o[2]);
			o[3] = bsdf->pdf(bRec);
		}
		{
			BSDFQueryRecord bRec(its, wi, Vector(0.0f));
			Float pdf = 0;
			Spectrum f = bsdf->sample(bRec, pdf, s);
			if (!f.isZero()) {
				o[4] = bRec.wo.x; o[5] = bRec.wo.y; o[6] = bRec.wo.z;
				f.toLinearRGB(o[7], o[8], o[9]);
				o[10] = pdf;
				o[11] = (float) bRec.sampledType;
			}
		}
		{
			BSDFQueryRecord bRec(its, wi, Vector(0.0f));
			Spectrum f = bsdf->sample(bRec, s);
			f.toLinearRGB(o[12], o[13], o[14]);
		}
	}
	sio::save(argv[4], out.data(), out.size() * 4);
	printf("{\"cmd\": \"bsdf\", \"queries\": %zu}\n", n);
	return 0;
}

/* queries: [p.xyz, s.x, s.y] -> 16 floats: sRec.p[3] sRec.n[3] d[3] pdf value[3] lumIndex pdfLuminaire pad
 * (Scene::sampleLuminaire without the visibility test, scene.cpp:396-415; pdfLuminaire 381-394) */
static int cmdEmitter(int argc, char **argv) {
	if (argc < 5) { std::cerr << "emitter SCENE QUERIES OUT\n"; return 1; }
	sio::Scene sf; sio::load(argv[2], sf);
	std::vector<float> q = sio::loadFloats(argv[3]);
	const size_t n = q.size() / 5;
	Built b = buildScene(sf, "");
	const std::vector<Luminaire *> &lums = b.scene->getLuminaires();
	std::vector<float> out(n * 16, 0.0f);
	for (size_t i = 0; i < n; ++i) {
		const float *e = &q[i * 5];
		Point p(e[0], e[1], e[2]);
		LuminaireSamplingRecord lRec;
		float *o = &out[i * 16];
		bool ok = b.scene->sampleLuminaire(p, 0.0f, lRec, Point2(e[3], e[4]), false);
		o[0] = lRec.sRec.p.x; o[1] = lRec.sRec.p.y; o[2] = lRec.sRec.p.z;
		o[3] = lRec.sRec.n.x; o[4] = lRec.sRec.n.y; o[5] = lRec.sRec.n.z;
		if (ok) {
			o[6] = lRec.d.x; o[7] = lRec.d.y; o[8] = lRec.d.z;
			o[9] = lRec.pdf;
			lRec.value.toLinearRGB(o[10], o[11], o[12]);
			size_t idx = 0;
			for (; idx < lums.size(); ++idx) if (lums[idx] == lRec.luminaire) break;
			o[13] = (float) idx;
			o[14] = b.scene->pdfLuminaire(p, lRec);
		} else {
			o[13] = -1.0f;
		}
	}
	sio::save(argv[4], out.data(), out.size() * 4);
	printf("{\"cmd\": \"emitter\", \"queries\": %zu, \"luminaires\": %zu}\n", n, lums.size());
	return 0;
}

/* samples: [x, y, lens.x, lens.y] raster positions -> rays [o.xyz mint d.xyz maxt] (perspective.cpp:77-112) */
static int cmdCamera(int argc, char **argv) {
	if (argc < 5) { std::cerr << "camera SCENE SAMPLES OUT\n"; return 1; }
	sio::Scene sf; sio::load(argv[2], sf);
	std::vector<float> q = sio::loadFloats(argv[3]);
	const size_t n = q.size() / 4;
	Built b = buildScene(sf, "", false);
	const Camera *cam = b.scene->getCamera();
	std::vector<float> out(n * 8);
	for (size_t i = 0; i < n; ++i) {
		Ray ray;
		cam->generateRay(Point2(q[i*4], q[i*4+1]), Point2(q[i*4+2], q[i*4+3]), 0.0f, ray);
		float *o = &out[i * 8];
		o[0] = ray.o.x; o[1] = ray.o.y; o[2] = ray.o.z; o[3] = ray.mint;
		o[4] = ray.d.x; o[5] = ray.d.y; o[6] = ray.d.z; o[7] = ray.maxt;
	}
	sio::save(argv[4], out.data(), out.size() * 4);
	printf("{\"cmd\": \"camera\", \"samples\": %zu}\n", n);
	return 0;
}

/* samples: [x, y, r, g, b, alpha] -> film accumulators [cropH][cropW][5], via a full-frame
 * ImageBlock with the filter's border (imageblock.h:80-138, renderproc.cpp:142-144) */
static int cmdSplat(int argc, char **argv) {
	if (argc < 5) { std::cerr << "splat SCENE SAMPLES OUT\n"; return 1; }
	sio::Scene sf; sio::load(argv[2], sf);
	std::vector<float> q = sio::loadFloats(argv[3]);
	const size_t n = q.size() / 6;
	Built b = buildScene(sf, "", false);
	Thread::getThread()->getLogger()->setLogLevel(EError);   /* silence the per-sample 'Invalid sample value' warnings */
	const TabulatedFilter *filter = b.film->getTabulatedFilter();
	int border = (int) std::ceil(std::max(filter->getFilterSize().x, filter->getFilterSize().y) - 0.5f);
	ref<ImageBlock> block = new ImageBlock(b.film->getCropSize(), border, true, true, false, false);
	block->setOffset(b.film->getCropOffset());
	block->setSize(b.film->getCropSize());
	block->clear();
	for (size_t i = 0; i < n; ++i) {
		const float *e = &q[i * 6];
		Spectrum s; s.fromLinearRGB(e[2], e[3], e[4]);
		block->putSample(Point2(e[0], e[1]), s, e[5], filter);   /* complain=true: invalid spectra are dropped, as in renderBlock */
	}
	b.film->clear();
	b.film->putImageBlock(block);
	sio::save(argv[4], b.film->raw().data(), b.film->raw().size() * 4);
	float table[256];
	for (int y = 0; y < 16; ++y) for (int x = 0; x < 16; ++x) table[y*16+x] = filter->lookup(x, y);
	if (argc > 5) sio::save(argv[5], table, sizeof(table));
	printf("{\"cmd\": \"splat\", \"samples\": %zu, \"border\": %d, \"filter_size\": [%g, %g]}\n", n, border,
		filter->getFilterSize().x, filter->getFilterSize().y);
	return 0;
}

/* Counts what a GUI or the `-r` flush thread would see: refresh signals (renderqueue.h:109-112) and finished blocks. */
class CountingListener : public RenderListener {
public:
	int refreshes, blocks; double weightAtLastRefresh; const RawFilm *film;
	CountingListener(const RawFilm *f) : refreshes(0), blocks(0), weightAtLastRefresh(0), film(f) { }
	void workBeginEvent(const RenderJob *, const RectangularWorkUnit *, int) { }
	void workEndEvent(const RenderJob *, const ImageBlock *) { blocks++; }
	void refreshEvent(const RenderJob *, const Bitmap *) {
		refreshes++;
		double w = 0;
		const std::vector<float> &raw = film->raw();
		for (size_t i = 4; i < raw.size(); i += 5) w += raw[i];
		weightAtLastRefresh = w;
	}
	void finishJobEvent(const RenderJob *, bool) { }
};

static int cmdRender(int argc, char **argv) {
	if (argc < 4) { std::cerr << "render SCENE OUT [threads] [integratorPlugin|-] [repeat]\n"; return 1; }
	sio::Scene sf; sio::load(argv[2], sf);
	int threads = argc > 4 ? atoi(argv[4]) : 1;
	std::string integ = argc > 5 ? argv[5] : "";
	if (integ == "-") integ = "";
	const int repeat = argc > 6 ? std::max(atoi(argv[6]), 1) : 1;     /* timing runs: the film then holds all passes */
	Scheduler *sched = Scheduler::getInstance();
	for (int i = 0; i < threads; ++i)
		sched->registerWorker(new LocalWorker(formatString("wrk%i", i)));
	sched->start();

	Built b = buildScene(sf, integ);
	ref<RenderQueue> queue = new RenderQueue();
	ref<CountingListener> listener = new CountingListener(b.film.get());
	queue->registerListener(listener);
	ref<RenderJob> job;
	double dt = 1e30; std::string all;
	for (int r = 0; r < repeat; ++r) {
		job = new RenderJob(formatString("ren%i", r), b.scene, queue, NULL, -1, -1, -1, false, false);
		double t0 = now();
		job->start();
		queue->waitLeft(0);
		const double d = now() - t0;
		job->join();
		dt = std::min(dt, d);
		all += (r ? ", " : "") + formatString("%.6f", d);
	}
	sio::save(argv[3], b.film->raw().data(), b.film->raw().size() * 4);
	std::string stats = Statistics::getInstance()->getStats();
	double samples = (double) sf.cropW * sf.cropH * sf.spp;
	printf("{\"cmd\": \"render\", \"integrator\": \"%s\", \"threads\": %d, \"width\": %u, \"height\": %u, \"spp\": %u, "
		"\"kd_build_s\": %.6f, \"render_s\": %.6f, \"render_s_all\": [%s], \"repeat\": %d, \"samples_per_s\": %.3f, "
		"\"refresh_events\": %d, \"film_weight_at_last_refresh\": %.3f, \"normal_rays\": \"%s\", \"shadow_rays\": \"%s\"}\n",
		integ.empty() ? (sf.mode == 0 ? "path" : "direct") : integ.c_str(), threads, sf.cropW, sf.cropH, sf.spp,
		b.buildSeconds, dt, all.c_str(), repeat, samples / dt, listener->refreshes, listener->weightAtLastRefresh,
		statValue(stats, "Normal rays traced").c_str(), statValue(stats, "Shadow rays traced").c_str());
	fflush(stdout);
	job = NULL; queue = NULL;
	sched->stop();
	return 0;
}

/* Load one mesh of a `.serialized` file through the reference's `serialized` shape plugin
 * (src/shapes/serialized.cpp) and dump what it hands to the kd-tree after TriMesh::configure. */
static int cmdMesh(int argc, char **argv) {
	if (argc < 5) { std::cerr << "mesh FILE SHAPEINDEX OUT [faceNormals flipNormals [16 x toWorld]]\n"; return 1; }
	Properties p("serialized");
	p.setString("filename", argv[2]);
	p.setInteger("shapeIndex", atoi(argv[3]));
	if (argc > 6) { p.setBoolean("faceNormals", atoi(argv[5]) != 0); p.setBoolean("flipNormals", atoi(argv[6]) != 0); }
	if (argc >= 23) {
		Matrix4x4 m;
		for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) m.m[i][j] = (Float) atof(argv[7 + 4 * i + j]);
		p.setTransform("toWorld", Transform(m));
	}
	ref<TriMesh> mesh = static_cast<TriMesh *>(PluginManager::getInstance()->createObject(MTS_CLASS(Shape), p));
	mesh->configure();
	FILE *f = fopen(argv[4], "wb");
	if (!f) { std::cerr << "cannot write output\n"; return 1; }
	const uint32_t hdr[4] = { (uint32_t) mesh->getVertexCount(), (uint32_t) mesh->getTriangleCount(),
		mesh->hasVertexNormals() ? 1u : 0u, mesh->hasVertexTexcoords() ? 1u : 0u };
	fwrite(hdr, 4, 4, f);
	fwrite(mesh->getVertexPositions(), sizeof(Point), hdr[0], f);
	if (hdr[2]) fwrite(mesh->getVertexNormals(), sizeof(Normal), hdr[0], f);
	if (hdr[3]) fwrite(mesh->getVertexTexcoords(), sizeof(Point2), hdr[0], f);
	fwrite(mesh->getTriangles(), sizeof(Triangle), hdr[1], f);
	fclose(f);
	printf("{\"cmd\": \"mesh\", \"vertices\": %u, \"triangles\": %u, \"normals\": %u, \"texcoords\": %u}\n", hdr[0], hdr[1], hdr[2], hdr[3]);
	return 0;
}

/* Write every mesh of a scene file with the reference's own writer: TriMesh::serialize(Stream *)
 * (trimesh.cpp:748-790) per mesh, then the offset table as the converter does (converter.cpp:293-296). */
static int cmdMeshSave(int argc, char **argv) {
	if (argc < 4) { std::cerr << "meshsave SCENE OUTFILE\n"; return 1; }
	sio::Scene sf; sio::load(argv[2], sf);
	Built b = buildScene(sf, "");
	ref<FileStream> out = new FileStream(argv[3], FileStream::ETruncReadWrite);
	out->setByteOrder(Stream::ELittleEndian);
	std::vector<uint32_t> dict;
	const std::vector<Shape *> &shapes = b.scene->getShapes();
	for (size_t i = 0; i < shapes.size(); ++i) {
		dict.push_back((uint32_t) out->getPos());
		static_cast<TriMesh *>(shapes[i])->serialize(out);
	}
	for (size_t i = 0; i < dict.size(); ++i) out->writeUInt(dict[i]);
	out->writeUInt((uint32_t) dict.size());
	out->close();
	printf("{\"cmd\": \"meshsave\", \"meshes\": %zu}\n", dict.size());
	return 0;
}

/* Run the reference's MFilm (src/films/mfilm.cpp) on given accumulators {r, g, b, alpha, weight} per pixel:
 * putImageBlock, then develop() writes OUT.m exactly as a render would. */
static int cmdDevelop(int argc, char **argv) {
	if (argc < 7) { std::cerr << "develop ACCUM WIDTH HEIGHT OUT spectra(0|1)\n"; return 1; }
	std::vector<float> acc = sio::loadFloats(argv[2]);
	const int w = atoi(argv[3]), h = atoi(argv[4]);
	if (acc.size() != (size_t) w * h * 5) { std::cerr << "accumulator size mismatch\n"; return 1; }
	Properties fp("mfilm");
	fp.setInteger("width", w); fp.setInteger("height", h);
	fp.setBoolean("spectra", atoi(argv[6]) != 0);
	ref<Film> film = static_cast<Film *>(PluginManager::getInstance()->createObject(MTS_CLASS(Film), fp));
	film->configure();
	ref<ImageBlock> block = new ImageBlock(Vector2i(w, h), 0, true, true, false, false);
	block->setOffset(Point2i(0, 0));
	block->setSize(Vector2i(w, h));
	block->clear();
	for (size_t i = 0; i < (size_t) w * h; ++i) {
		Spectrum s; s.fromLinearRGB(acc[i * 5], acc[i * 5 + 1], acc[i * 5 + 2]);
		block->setPixel(i, s);
		block->setWeight(i, acc[i * 5 + 4]);
	}
	film->clear();
	film->putImageBlock(block);
	film->develop(argv[5]);
	printf("{\"cmd\": \"develop\", \"pixels\": %d}\n", w * h);
	return 0;
}

/* Closest hit -> interpolated texture coordinates (skdtree.h:408-415) and BSDF::getDiffuseReflectance(its), which for
 * lambertian / microfacet / phong is the (possibly textured) reflectance at the hit (times the diffuse amount). */
static int cmdUv(int argc, char **argv) {
	if (argc < 5) { std::cerr << "uv SCENE RAYS OUT\n"; return 1; }
	sio::Scene sf; sio::load(argv[2], sf);
	std::vector<float> rays = sio::loadFloats(argv[3]);
	const size_t n = rays.size() / 8;
	Built b = buildScene(sf, "");
	std::vector<float> out(n * 5, 0.0f);
	for (size_t i = 0; i < n; ++i) {
		const float *e = &rays[i * 8];
		Ray ray(Point(e[0], e[1], e[2]), Vector(e[4], e[5], e[6]), e[3], e[7], 0.0f);
		Intersection its;
		if (!b.scene->rayIntersect(ray, its)) continue;
		out[i * 5] = its.uv.x; out[i * 5 + 1] = its.uv.y;
		if (its.shape->hasBSDF())
			its.shape->getBSDF()->getDiffuseReflectance(its).toLinearRGB(out[i * 5 + 2], out[i * 5 + 3], out[i * 5 + 4]);
	}
	sio::save(argv[4], out.data(), out.size() * 4);
	printf("{\"cmd\": \"uv\", \"rays\": %zu}\n", n);
	return 0;
}

/* samples: [x, y, lens.x, lens.y] -> 12 floats: its.uv, dudx, dudy, dvdx, dvdy, diffuse and specular reflectance at
 * the hit of the camera's ray differential (integrator.cpp:163 -> path.cpp:70: its.getBSDF(ray) runs
 * Intersection::computePartials when a texture of the BSDF filters) */
static int cmdTexLookup(int argc, char **argv) {
	if (argc < 5) { std::cerr << "texlookup SCENE SAMPLES OUT\n"; return 1; }
	sio::Scene sf; sio::load(argv[2], sf);
	std::vector<float> q = sio::loadFloats(argv[3]);
	const size_t n = q.size() / 4;
	Built b = buildScene(sf, "");
	const Camera *camera = b.scene->getCamera();
	std::vector<float> out(n * 12, 0.0f);
	for (size_t i = 0; i < n; ++i) {
		const float *e = &q[i * 4];
		RayDifferential ray;
		camera->generateRayDifferential(Point2(e[0], e[1]), Point2(e[2], e[3]), 0.0f, ray);
		ray.scaleDifferential(1.0f / std::sqrt((Float) sf.spp));        /* SampleIntegrator::renderBlock, integrator.cpp:145,165 */
		Intersection its;
		if (!b.scene->rayIntersect(ray, its)) continue;
		float *o = &out[i * 12];
		o[0] = its.uv.x; o[1] = its.uv.y;
		const BSDF *bsdf = its.getBSDF(ray);
		if (!bsdf) continue;
		if (its.hasUVPartials) { o[2] = its.dudx; o[3] = its.dudy; o[4] = its.dvdx; o[5] = its.dvdy; }
		bsdf->getDiffuseReflectance(its).toLinearRGB(o[6], o[7], o[8]);
	}
	sio::save(argv[4], out.data(), out.size() * 4);
	printf("{\"cmd\": \"texlookup\", \"samples\": %zu}\n", n);
	return 0;
}

static int dispatch(int argc, char **argv) {
	if (argc < 2) { std::cerr << "usage: mts_oracle trace|triaccel|bsdf|emitter|camera|splat|render|mesh|meshsave|develop|uv|texlookup ...\n"; return 1; }
	std::string cmd = argv[1];
	if (cmd == "trace") return cmdTraceImpl(argc, argv);
	if (cmd == "triaccel") return cmdTriAccel(argc, argv);
	if (cmd == "bsdf") return cmdBsdf(argc, argv);
	if (cmd == "emitter") return cmdEmitter(argc, argv);
	if (cmd == "camera") return cmdCamera(argc, argv);
	if (cmd == "splat") return cmdSplat(argc, argv);
	if (cmd == "render") return cmdRender(argc, argv);
	if (cmd == "mesh") return cmdMesh(argc, argv);
	if (cmd == "meshsave") return cmdMeshSave(argc, argv);
	if (cmd == "develop") return cmdDevelop(argc, argv);
	if (cmd == "uv") return cmdUv(argc, argv);
	if (cmd == "texlookup") return cmdTexLookup(argc, argv);
	std::cerr << "unknown command " << cmd << "\n";
	return 1;
}
