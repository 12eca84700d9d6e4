ssh test@127.0.0.1 "kill -9 3085"
ssh test@127.0.0.1 "kill -9 3086"
ssh test@127.0.0.1 "kill -9 3087"
ssh test@127.0.0.1 "kill -9 3088"
kill -9 2946
