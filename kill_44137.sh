ssh test@127.0.0.1 "kill -9 3275"
ssh test@127.0.0.1 "kill -9 3276"
ssh test@127.0.0.1 "kill -9 3277"
kill -9 2946
