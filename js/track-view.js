// Same module path and default export as lib/integration/mapping/track-view.js
module.exports = require('./mapping').TrackView;
module.exports.default = module.exports;
